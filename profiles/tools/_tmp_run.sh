timeout 300 python -m pytest tests/test_gpu_mlp.py -m gpu -x -q 2>&1 | tail -4
for m in on torch; do
python bench.py --mlps $m --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2i_bench_mlps_$m.json 2> gpurun_out/r2i_bench_mlps_$m.err; echo "== mlps $m rc=$?"; tail -c 400 gpurun_out/r2i_bench_mlps_$m.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2i_bench_mlps_$m.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"])
print({k:(v["ms_per_launch"],v["launches"]) for k,v in d["kernels"].items()})
print(d.get("reference_cuda_b200"))
PY
done
