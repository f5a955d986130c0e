#!/bin/bash
# Round-2 closing run on one GPU: whole GPU suite in one process, smoke(), the default bench line,
# the launch list of the bench and the --set full capture of the fused kernel (pair mode), and the
# GAE mode sweep. (Reference arm, H1 / GAE / interval captures: scripts/gpu_r2_final.sh, unchanged.)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x -p no:cacheprovider > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_final.log 2> gpurun_out/bench_final.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench_final.log
python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-extras > gpurun_out/plain_bench.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-extras > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
python scripts/prof_h1.py 2048 fused > gpurun_out/plain_fused.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:grpo_fused -s 1 -c 2 -f -o gpurun_out/prof_fused \
    python scripts/prof_h1.py 2048 fused > gpurun_out/ncu_fused.log 2>&1
echo "fused rc=$?"; tail -2 gpurun_out/ncu_fused.log
timeout 600 python scripts/time_gae_modes.py > gpurun_out/time_gae_modes_full.log 2>&1; echo "gae modes rc=$?"; tail -3 gpurun_out/time_gae_modes_full.log
