#!/bin/bash
mkdir -p gpurun_out
python scripts/prof_h1.py 0 gae_team > gpurun_out/plain_gae_team.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gae_serial_team -s 1 -c 1 -f -o gpurun_out/prof_gae_team \
    python scripts/prof_h1.py 0 gae_team > gpurun_out/ncu_gae_team.log 2>&1
echo "team rc=$?"; tail -2 gpurun_out/ncu_gae_team.log
timeout 900 python scripts/time_misc.py > gpurun_out/kernels_vs_reference.txt 2>&1; echo "misc rc=$?"; tail -12 gpurun_out/kernels_vs_reference.txt
