#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
N=8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29701 bench.py --gpus $N --steps 30 > gpurun_out/z_C0_n8.json 2> gpurun_out/z_C0_n8.err; echo "C0 rc=$?"
timeout 300 $TR --master-port 29702 bench.py --gpus $N --workload C4 --views-per-step 8 --steps 10 > gpurun_out/z_C4v8_n8.json 2> gpurun_out/z_C4v8_n8.err; echo "C4 rc=$?"
timeout 300 $TR --master-port 29703 bench.py --gpus $N --workload C5 --mode sequence --steps 2 > gpurun_out/z_C5seq_n8.json 2> gpurun_out/z_C5seq_n8.err; echo "C5 rc=$?"
timeout 300 $TR --master-port 29704 tests/multi_gpu_check.py > gpurun_out/z_check_n8.log 2>&1; echo "check rc=$?"
for f in z_C0_n8 z_C4v8_n8 z_C5seq_n8; do python -c "
import json
d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1]); print('$f', d['ms_per_step'], d['value'], d['e2e']['value'], d.get('allreduce_checked'))
"; done
grep view_split gpurun_out/z_check_n8.log | cut -c1-300
