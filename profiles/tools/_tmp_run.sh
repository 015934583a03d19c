#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
for wl in C4 C5; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --mode split --workload $wl --steps 16 > gpurun_out/split_${wl}_n2.json 2> gpurun_out/split_${wl}_n2.err; echo "bench $wl rc=$?"
grep -i "disagree" gpurun_out/split_${wl}_n2.err | head -n 5
done
python bench.py --mode split --workload C5 --steps 16 > gpurun_out/split_C5_n1.json 2> gpurun_out/split_C5_n1.err
