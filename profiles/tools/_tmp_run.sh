N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
run() { tag=$1; shift; env "$@" $TR bench.py --gpus $N --steps 30 --warmup 5 $EXTRA > gpurun_out/r2f_bench_${tag}_n$N.json 2> gpurun_out/r2f_bench_${tag}_n$N.err; echo "== $tag rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2f_bench_${tag}_n$N.json").read().strip().splitlines()[-1])
    print(d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("allreduce_checked"), d.get("per_rank_ms_per_step"), d.get("per_rank_ms_per_frame"))
    print({k:v["ms_per_launch"] for k,v in d["kernels"].items() if "allred" in k or "gauss" in k})
except Exception as e:
    print("no json", e); print(open("gpurun_out/r2f_bench_${tag}_n$N.err").read()[-1500:])
PY
}
EXTRA=""
run C0_default A=1
run C0_ov16x4 RG_DP_OVERLAP=1 RG_DP_BLOCKS=16 RG_DP_CHUNKS=4
run C0_ov32x8 RG_DP_OVERLAP=1 RG_DP_BLOCKS=32 RG_DP_CHUNKS=8
run C0_mm_ov16x4 RG_ALLREDUCE=multimem RG_DP_OVERLAP=1 RG_DP_BLOCKS=16 RG_DP_CHUNKS=4
EXTRA="--workload C5 --mode sequence --steps 3"
run C5_seq A=1
