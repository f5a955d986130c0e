#!/bin/bash
# Round-2 evidence on one GPU: default bench line, reference arm, kernels vs the reference's own CUDA
# kernels + host-overhead timings, ncu launch list and --set full captures (each only after the
# same command has exited 0 without ncu).
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_final.log 2> gpurun_out/bench_final.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref_final.log 2>&1; echo "ref rc=$?"
timeout 900 python scripts/time_misc.py > gpurun_out/kernels_vs_reference.txt 2>&1; echo "misc rc=$?"; tail -8 gpurun_out/kernels_vs_reference.txt
bash scripts/gpu_prof.sh
bash scripts/gpu_prof_interval.sh
