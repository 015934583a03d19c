python -m pytest tests -m gpu -q 2>&1 | tail -8
for w in "C3 --views-per-step 4" "C4 --views-per-step 8 --steps 10"; do n=$(echo $w | cut -d" " -f1); python bench.py --workload $w --no-cpu-baseline --no-ref-cuda > gpurun_out/r2c_bench_$n.json 2> gpurun_out/r2c_bench_$n.err; echo "== $w rc=$?"; tail -c 300 gpurun_out/r2c_bench_$n.err; head -c 700 gpurun_out/r2c_bench_$n.json; echo; done
