#!/bin/bash
# Runs on the GPU box (under gpurun): smoke, each GPU test file in its own process, short bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke" ; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
tail -3 gpurun_out/smoke.log
for f in tests/test_*_gpu.py tests/test_zz_no_alloc.py; do
  n=$(basename $f .py)
  timeout 900 python -m pytest $f -q -m gpu --maxfail=25 -p no:cacheprovider > gpurun_out/$n.log 2>&1
  echo "== $n rc=$?"; tail -4 gpurun_out/$n.log
done
echo "== bench"; timeout 900 python bench.py --steps 2 --warmup 1 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
