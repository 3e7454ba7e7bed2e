#!/bin/bash
# compute-sanitizer over small shapes of the tcgen05 kernels (SURVEY.md section 5 pin).  ONE tool per call (and per gpurun
# call): tools/sanitize.sh memcheck|racecheck|synccheck [pytest -k expression]
#   gpurun --timeout 1500 -- 'tools/sanitize.sh memcheck > gpurun_out/sanitize_memcheck.log 2>&1'
# The summary lines (ERROR SUMMARY / RACECHECK SUMMARY) are what profiles/ keeps.
set -u
tool=${1:-memcheck}
expr=${2:-"grad_vs_oracle and (128 or 1000) or weight_gradient_stage or forward_recurrence_stage and 37 or backward_recurrence_stage and 37"}
cd "$(dirname "$0")/.."
timeout 1200 compute-sanitizer --tool "$tool" --print-limit 20 \
  python -m pytest tests/test_mlp_tc_gpu.py tests/test_mlp_tc64_gpu.py tests/test_lstm_tc_gpu.py -m gpu -q -x -k "$expr" 2>&1 | tail -60
