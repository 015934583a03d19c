python -m pytest tests/test_gpu_fused.py tests/test_gpu_fullsize.py tests/test_gpu_raster.py -m gpu -q -x 2>&1 | tail -4
python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-ref-cuda > gpurun_out/r2n_bench.json 2> gpurun_out/r2n_bench.err; echo rc=$?
python - <<PY
import json
d=json.loads(open("gpurun_out/r2n_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"])
print({k:v["ms_per_launch"] for k,v in d["kernels"].items()})
PY
