#!/bin/bash
# End-of-round evidence: default bench line, reference arm, launch list and ncu captures.
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_final.log 2> gpurun_out/bench_final.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_final.log 2>&1; echo "ref rc=$?"
bash scripts/gpu_prof.sh
bash scripts/gpu_prof_interval.sh
