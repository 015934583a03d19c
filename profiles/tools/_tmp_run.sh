N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
$TR tests/multi_gpu_check.py > gpurun_out/r2e_allreduce_check_n$N.log 2> gpurun_out/r2e_allreduce_check_n$N.err; echo "check rc=$?"; grep -c true gpurun_out/r2e_allreduce_check_n$N.log; grep -i "false\|skipped" gpurun_out/r2e_allreduce_check_n$N.log
run() { tag=$1; shift; env "$@" $TR bench.py --gpus $N --steps 30 --warmup 5 $EXTRA > gpurun_out/r2e_bench_${tag}_n$N.json 2> gpurun_out/r2e_bench_${tag}_n$N.err; echo "== $tag rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2e_bench_${tag}_n$N.json").read().strip().splitlines()[-1])
    print(d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("allreduce_checked"), d.get("nvlink"))
except Exception as e:
    print("no json", e); print(open("gpurun_out/r2e_bench_${tag}_n$N.err").read()[-800:])
PY
}
EXTRA=""
run C0_default A=1
run C0_ov16x4 RG_DP_OVERLAP=1 RG_DP_BLOCKS=16 RG_DP_CHUNKS=4
run C0_ov32x8 RG_DP_OVERLAP=1 RG_DP_BLOCKS=32 RG_DP_CHUNKS=8
run C0_ar148 RG_AR_BLOCKS=148
run C0_ar592 RG_AR_BLOCKS=592
run C0_p2p RG_ALLREDUCE=p2p
EXTRA="--workload C4 --views-per-step 8 --steps 10"
run C4_v8 A=1
EXTRA="--workload C5 --mode sequence --steps 3"
run C5_seq A=1
