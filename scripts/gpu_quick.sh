#!/bin/bash
# quick loop: H1 + loss tests, micro timing of the H1 kernels, short bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_logprob_gpu.py tests/test_loss_advnorm_gpu.py -q -m gpu -x -p no:cacheprovider > gpurun_out/quick_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/quick_tests.log
timeout 300 python scripts/time_h1.py > gpurun_out/time_h1.log 2>&1; echo "time rc=$?"; cat gpurun_out/time_h1.log | tail -12
timeout 600 python bench.py --steps 5 --warmup 2 --no-e2e --no-cpu --no-extras > gpurun_out/bench_quick.log 2>gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/bench_quick.log').read().strip().splitlines()[-1])
    r=d['roofline']
    print('value %.3fM tok/s  ms/step %.2f  bwd %.0f GB/s (%.3f)  fwd %.0f GB/s (%.3f)  both %.3f clocks %s' % (d['value']/1e6, d['ms_per_step'], r['achieved'], r['frac'], r['fwd']['achieved'], r['fwd']['frac'], r['fwd_bwd_frac_of_peak'], d['clocks']))
except Exception as e:
    print('bench parse failed', e); print(open('gpurun_out/bench_quick.err').read()[-2000:])
PY
