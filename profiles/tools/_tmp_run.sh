timeout 120 python -m pytest tests/test_gpu_mlp.py -m gpu -x -q 2>&1 | tail -25
cat gpurun_out/mlp_fwd_*.json 2>/dev/null
