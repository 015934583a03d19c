#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
python -m pytest tests/test_gpu_raster.py tests/test_gpu_fused.py -x -q -m gpu > gpurun_out/band_tests.log 2>&1; echo "tests rc=$?"
tail -n 3 gpurun_out/band_tests.log
run() { name=$1; shift
  python bench.py "$@" --no-cpu-baseline --no-ref-cuda > gpurun_out/$name.json 2> gpurun_out/$name.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/$name.json').read().strip().splitlines()[-1]); print('$name', d['ms_per_step'], d['value'], {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items() if k.startswith('tile')})
"
}
run b2_c4 --workload C4 --views-per-step 8 --steps 4
run b2_c5 --workload C5 --mode sequence --steps 2
RG_BIN_MAXROWS=4096 run b2_c5_r4096 --workload C5 --mode sequence --steps 2
run b2_c3 --workload C3 --views-per-step 4 --steps 10
run b2_c0 --steps 30
