#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
python bench.py --steps 30 --baselines > gpurun_out/final_n1.json 2> gpurun_out/final_n1.err; echo "c0 rc=$?"
python bench.py --mlps on --steps 10 > gpurun_out/final_mlps_on.json 2> gpurun_out/final_mlps_on.err; echo "mlps rc=$?"
python bench.py --workload C2 --steps 30 --no-cpu-baseline --no-ref-cuda > gpurun_out/final_C2.json 2> gpurun_out/final_C2.err
python bench.py --workload C3 --views-per-step 4 --steps 10 --no-cpu-baseline --no-ref-cuda > gpurun_out/final_C3v4.json 2> gpurun_out/final_C3v4.err
python bench.py --workload C4 --views-per-step 8 --steps 6 --no-cpu-baseline --no-ref-cuda > gpurun_out/final_C4v8.json 2> gpurun_out/final_C4v8.err
python bench.py --workload C5 --mode sequence --steps 2 > gpurun_out/final_C5seq.json 2> gpurun_out/final_C5seq.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_reference_arm.json 2> gpurun_out/final_reference_arm.err
ncu --set full --clock-control none -k regex:mlp_ -c 6 --csv --page raw --log-file gpurun_out/mlp_full_raw.csv python profiles/tools/mlp_time.py 300000 1 > gpurun_out/ncu_mlp2.log 2>&1; echo "ncu rc=$?"
for f in final_n1 final_mlps_on final_C2 final_C3v4 final_C4v8 final_C5seq final_reference_arm; do python -c "
import json,sys
d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1]); print('$f', d.get('ms_per_step'), d.get('value'), d.get('e2e',{}).get('value'))
"; done
