#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_mlp.py -x -q -m gpu > gpurun_out/mlp_tests.log 2>&1; echo "tests rc=$?"
tail -n 3 gpurun_out/mlp_tests.log
timeout 120 python profiles/tools/mlp_time.py 300000 5
timeout 300 python bench.py --mlps on --steps 10 --no-cpu-baseline --no-ref-cuda > gpurun_out/mlp_on.json 2> gpurun_out/mlp_on.err
python -c "
import json,sys
d=json.loads(open('gpurun_out/mlp_on.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items() if k.startswith('mlp')})
"
