#!/bin/bash
# sustained A/B of the fused step through bench.py (20 steps), pair mode off / on, twice
mkdir -p gpurun_out
O=gpurun_out/pair_bench.log
: > $O
for rep in 1 2; do
for m in 0 1; do
  B200RL_FUSED_PAIR=$m timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu --no-extras 2>/dev/null | grep '^{"metric"' | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('pair=$m', 'value', d['value'], 'fused', json.dumps(d.get('fused_step')), 'clocks', d.get('clocks'))
" >> $O 2>&1
done
done
for d in 6000 8000 10000 12000; do
echo "--- pair delay $d" >> $O; B200RL_FUSED_PAIR=1 B200RL_FUSED_PAIR_DELAY_NS=$d timeout 120 python scripts/time_fused.py 40 >> $O 2>&1
done
cat $O
