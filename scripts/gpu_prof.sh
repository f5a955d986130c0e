#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-extras > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-extras > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
python scripts/prof_h1.py 2048 h1 > gpurun_out/plain_h1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:logprob_ -s 4 -c 4 -o gpurun_out/prof_h1 \
    python scripts/prof_h1.py 2048 h1 > gpurun_out/ncu_h1.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu_h1.log
python scripts/prof_h1.py 2048 fused > gpurun_out/plain_fused.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:grpo_fused -s 1 -c 2 -o gpurun_out/prof_fused \
    python scripts/prof_h1.py 2048 fused > gpurun_out/ncu_fused.log 2>&1
echo "fused rc=$?"; tail -2 gpurun_out/ncu_fused.log
python scripts/prof_h1.py 0 gae > gpurun_out/plain_gae.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gae_ -s 3 -c 3 -o gpurun_out/prof_gae \
    python scripts/prof_h1.py 0 gae > gpurun_out/ncu_gae.log 2>&1
echo "gae rc=$?"; tail -2 gpurun_out/ncu_gae.log
