#!/bin/bash
for lib in libb200rl tune_w20_s2 tune_w24_s2 tune_w12_s4; do
  export B200RL_LIB=$PWD/areal_vllm_b200/$lib.so
  echo "=== $lib"
  timeout 300 python scripts/time_h1.py 2>&1 | grep -E "^fwd|^bwd" | head -5
  timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu --no-extras 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('  bench: %.3fM tok/s  %.2f ms/step  bwd %.0f (%.3f)  fwd %.0f (%.3f)  both %.3f  clk %s %s' % (d['value']/1e6, d['ms_per_step'], r['achieved'], r['frac'], r['fwd']['achieved'], r['fwd']['frac'], r['fwd_bwd_frac_of_peak'], d['clocks']['sm_mhz'], d['clocks']['reasons']))"
done
unset B200RL_LIB
timeout 600 python -m pytest tests/test_gae_gpu.py tests/test_reference_cuda_gpu.py -q -m gpu -x -p no:cacheprovider 2>&1 | tail -2
timeout 600 python scripts/time_misc.py 2>&1 | grep -E "GAE" | grep serial
