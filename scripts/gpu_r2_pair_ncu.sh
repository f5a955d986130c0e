#!/bin/bash
mkdir -p gpurun_out
B200RL_FUSED_PAIR=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:grpo_fused -s 1 -c 1 -f -o gpurun_out/prof_fused_pair \
    python scripts/prof_h1.py 2048 fused > gpurun_out/ncu_fused_pair.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_fused_pair.log
