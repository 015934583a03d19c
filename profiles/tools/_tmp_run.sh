#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
python -m pytest tests/test_gpu_raster.py tests/test_gpu_fused.py tests/test_gpu_fullsize.py -x -q -m gpu > gpurun_out/band_tests.log 2>&1; echo "tests rc=$?"
tail -n 5 gpurun_out/band_tests.log
python bench.py --workload C4 --views-per-step 8 --steps 6 --no-cpu-baseline --no-ref-cuda > gpurun_out/band_C4v8.json 2> gpurun_out/band_C4v8.err
python bench.py --workload C3 --views-per-step 4 --steps 10 --no-cpu-baseline --no-ref-cuda > gpurun_out/band_C3v4.json 2> gpurun_out/band_C3v4.err
python bench.py --steps 30 --no-cpu-baseline --no-ref-cuda > gpurun_out/band_C0.json 2> gpurun_out/band_C0.err
for f in band_C4v8 band_C3v4 band_C0; do python -c "
import json,sys
d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1]); print('$f', d['ms_per_step'], d['value'], {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items() if k.startswith('tile')})
"; done
