N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
$TR tests/multi_gpu_check.py > gpurun_out/r2d_allreduce_check_n$N.log 2> gpurun_out/r2d_allreduce_check_n$N.err; echo "check rc=$?"; cat gpurun_out/r2d_allreduce_check_n$N.log; tail -5 gpurun_out/r2d_allreduce_check_n$N.err
$TR bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/r2d_bench_C0_n$N.json 2> gpurun_out/r2d_bench_C0_n$N.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r2d_bench_C0_n$N.err; python - <<PY
import json
d=json.loads(open("gpurun_out/r2d_bench_C0_n$N.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("allreduce_checked"), d.get("nvlink"), d["roofline"])
print({k:v["ms_per_launch"] for k,v in d["kernels"].items()})
PY
