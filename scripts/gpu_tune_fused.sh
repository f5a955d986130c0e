#!/bin/bash
timeout 600 python -m pytest tests/test_fused_gpu.py tests/test_canary_gpu.py -q -m gpu -x -p no:cacheprovider 2>&1 | tail -2
for lib in libb200rl tune_noskew; do
  export B200RL_LIB=$PWD/areal_vllm_b200/$lib.so
  echo "=== $lib"; timeout 300 python scripts/time_fused.py 30 2>&1 | tail -1; timeout 300 python scripts/time_fused.py 600 2>&1 | tail -1
done
unset B200RL_LIB
python scripts/prof_h1.py 2048 fused > gpurun_out/plain_fused.log 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:grpo_fused -s 1 -c 1 --csv --log-file gpurun_out/fused_dram.csv python scripts/prof_h1.py 2048 fused > /dev/null 2>&1
grep -E "dram__bytes|gpu__time|hit_rate" gpurun_out/fused_dram.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}'
