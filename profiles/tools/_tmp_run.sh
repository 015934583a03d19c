#!/bin/bash
cd /root/repo
N=2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
p=29720
for cfg in "4 296" "8 296" "8 592" "4 592" "2 592" "8 1184"; do set -- $cfg; p=$((p+1))
RG_AR_UNROLL=$1 RG_AR_BLOCKS=$2 timeout 120 $TR --master-port $p profiles/tools/ar_time.py 2>/dev/null | grep world
done
