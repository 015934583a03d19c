#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
python -m pytest tests/test_gpu_knn_and_api.py -x -q -m gpu -k "l1 or loss" > gpurun_out/t.log 2>&1; echo "tests rc=$?"; tail -n 2 gpurun_out/t.log
python bench.py --steps 30 --no-cpu-baseline --no-ref-cuda > gpurun_out/c0.json 2> gpurun_out/c0.err
python -c "
import json,sys
d=json.loads(open('gpurun_out/c0.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items() if k in ('l1_loss','fk_forward','fk_backward','tile_offsets')})
"
