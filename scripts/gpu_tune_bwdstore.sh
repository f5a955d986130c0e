#!/bin/bash
for lib in libb200rl bwd_ef; do
  export B200RL_LIB=$PWD/areal_vllm_b200/$lib.so
  echo "=== $lib"
  timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu --no-extras --no-fused 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('  bench: %.3fM tok/s  %.2f ms/step  bwd %.0f (%.3f)  fwd %.0f (%.3f)  both %.3f  clk %s %s' % (d['value']/1e6, d['ms_per_step'], r['achieved'], r['frac'], r['fwd']['achieved'], r['fwd']['frac'], r['fwd_bwd_frac_of_peak'], d['clocks']['sm_mhz'], d['clocks']['reasons']))"
done
