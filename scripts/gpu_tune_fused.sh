#!/bin/bash
for lib in libb200rl tune_fw20_s2 tune_fw24_s2 tune_fw12_s4; do
  export B200RL_LIB=$PWD/areal_vllm_b200/$lib.so
  echo "=== $lib"; timeout 300 python scripts/time_fused.py 30 2>&1 | tail -1; timeout 300 python scripts/time_fused.py 600 2>&1 | tail -1
done
