#!/bin/bash
for lib in iv_tm32768 iv_tm65536 iv_tm131072 iv_tm262144; do
  export B200RL_LIB=$PWD/areal_vllm_b200/$lib.so
  echo "=== $lib"
  timeout 300 python scripts/exp_interval.py 2>&1 | tail -6
done
unset B200RL_LIB
timeout 600 python -m pytest tests/test_intervals_gpu.py tests/test_move_intervals_gpu.py tests/test_reference_cuda_gpu.py tests/test_canary_gpu.py -q -m gpu -x -p no:cacheprovider 2>&1 | tail -3
