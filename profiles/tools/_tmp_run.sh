python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2m_bench_reference_arm.json 2> gpurun_out/r2m_ref.err; echo "ref rc=$?"
python bench.py --steps 30 --warmup 5 --baselines > gpurun_out/r2m_bench_n1.json 2> gpurun_out/r2m_n1.err; echo "ours rc=$?"; tail -c 300 gpurun_out/r2m_n1.err
python - <<PY
import json
for f in ("r2m_bench_reference_arm","r2m_bench_n1"):
    d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
    print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["config"]["workload"])
    for k in ("cpu_baseline","reference_cuda_b200","reference_cpu_skinning","reference_cuda_rasterizer","step_hbm","roofline","clocks"):
        if k in d: print("  ",k,d[k])
PY
