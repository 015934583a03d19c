timeout 300 python -m pytest tests/test_gpu_mlp.py -m gpu -x -q 2>&1 | tail -4
python bench.py --mlps on --steps 10 --warmup 3 --no-cpu-baseline --no-ref-cuda > gpurun_out/r2k_bench_mlps_on.json 2> gpurun_out/r2k_bench_mlps_on.err; echo "rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r2k_bench_mlps_on.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"])
print({k:(v["ms_per_launch"],v["launches"]) for k,v in d["kernels"].items() if "mlp" in k})
PY
