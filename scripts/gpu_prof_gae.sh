#!/bin/bash
mkdir -p gpurun_out
python scripts/prof_h1.py 0 gae > gpurun_out/plain_gae.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gae_ -s 3 -c 3 -o gpurun_out/prof_gae \
    python scripts/prof_h1.py 0 gae > gpurun_out/ncu_gae.log 2>&1
echo "gae rc=$?"; tail -2 gpurun_out/ncu_gae.log
