#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_mlp.py -x -q -m gpu > gpurun_out/mlp_tests.log 2>&1; echo "tests rc=$?"
tail -n 12 gpurun_out/mlp_tests.log
