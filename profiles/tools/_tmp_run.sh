#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
python -m pytest tests/test_gpu_fused.py tests/test_gpu_skin.py -x -q -m gpu > gpurun_out/cache_tests.log 2>&1; echo "tests rc=$?"
tail -n 3 gpurun_out/cache_tests.log
python bench.py --workload C5 --mode sequence --steps 2 > gpurun_out/cache_c5.json 2> gpurun_out/cache_c5.err
python -c "
import json,sys
d=json.loads(open('gpurun_out/cache_c5.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items()})
"
python bench.py --steps 30 --no-cpu-baseline --no-ref-cuda > gpurun_out/cache_c0.json 2> gpurun_out/cache_c0.err
python -c "
import json,sys
d=json.loads(open('gpurun_out/cache_c0.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items()})
"
