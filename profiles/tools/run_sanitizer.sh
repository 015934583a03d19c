#!/bin/bash
# compute-sanitizer runs of the fused forward with whole CTAs culled (the bulk-load-before-exit case) and of one
# small fused forward + backward; logs go to gpurun_out/ and are summarised under profiles/.
cd "$(dirname "$0")/../.." || exit 1
mkdir -p gpurun_out
T="tests/test_gpu_fused.py::test_fused_forward_with_whole_blocks_culled"
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 python -m pytest -x -q -m gpu "$T" \
      > gpurun_out/sanitizer_${tool}_culled.log 2>&1
  echo "$tool culled-blocks exit=$?" | tee -a gpurun_out/sanitizer_summary.txt
done
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest -x -q -m gpu \
    "tests/test_gpu_fused.py::test_fused_matches_unfused[ragged-over5]" > gpurun_out/sanitizer_memcheck_fused_ragged.log 2>&1
echo "memcheck fused fwd+bwd (ragged) exit=$?" | tee -a gpurun_out/sanitizer_summary.txt
