#!/bin/bash
mkdir -p gpurun_out
python scripts/prof_interval.py > gpurun_out/plain_interval.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:interval_copy\|interval_move -s 3 -c 3 -o gpurun_out/prof_interval \
    python scripts/prof_interval.py > gpurun_out/ncu_interval.log 2>&1
echo "interval rc=$?"; tail -2 gpurun_out/ncu_interval.log
