#!/bin/bash
# A/B of the fused GRPO kernel: one row per CTA vs. pair mode (clusters of two CTAs, two row streams)
mkdir -p gpurun_out
O=gpurun_out/pair.log
: > $O
B200RL_FUSED_PAIR=0 timeout 120 python scripts/fused_pair_check.py > gpurun_out/pair_chk0.txt 2>>$O
B200RL_FUSED_PAIR=1 B200RL_FUSED_PAIR_VERBOSE=1 timeout 120 python scripts/fused_pair_check.py > gpurun_out/pair_chk1.txt 2>>$O
echo "check rc=$?" >> $O
if cmp -s gpurun_out/pair_chk0.txt gpurun_out/pair_chk1.txt; then echo "IDENTICAL" >> $O; else echo "DIFFERENT" >> $O; diff gpurun_out/pair_chk0.txt gpurun_out/pair_chk1.txt >> $O; fi
B200RL_FUSED_PAIR=1 timeout 300 python -m pytest tests/test_fused_gpu.py -x -q -m gpu 2>&1 | tail -3 >> $O
echo "--- base" >> $O;  B200RL_FUSED_PAIR=0 timeout 120 python scripts/time_fused.py 40 >> $O 2>&1
for d in 0 4000 8000; do
echo "--- pair delay $d" >> $O; B200RL_FUSED_PAIR=1 B200RL_FUSED_PAIR_DELAY_NS=$d timeout 120 python scripts/time_fused.py 40 >> $O 2>&1
done
cat $O
