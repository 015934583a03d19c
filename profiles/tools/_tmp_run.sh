#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
python -m pytest tests -q -m gpu > gpurun_out/full_gpu_tests.log 2>&1; echo "tests rc=$?"
tail -n 4 gpurun_out/full_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
python bench.py > gpurun_out/default_bench.json 2> gpurun_out/default_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/default_bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'], d['gpu_launches'], d['steps'])
"
