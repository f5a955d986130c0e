#!/bin/bash
# Fused GRPO kernel, one row per CTA (B200RL_FUSED_PAIR=0) vs pair mode: outputs must be identical,
# then burst timings of both.
mkdir -p gpurun_out
O=gpurun_out/pair.log
: > $O
B200RL_FUSED_PAIR=0 timeout 120 python scripts/fused_pair_check.py > gpurun_out/pair_chk0.txt 2>>$O
B200RL_FUSED_PAIR=1 B200RL_FUSED_PAIR_VERBOSE=1 timeout 120 python scripts/fused_pair_check.py > gpurun_out/pair_chk1.txt 2>>$O
if cmp -s gpurun_out/pair_chk0.txt gpurun_out/pair_chk1.txt; then echo "IDENTICAL" >> $O; else echo "DIFFERENT" >> $O; fi
for m in 0 1; do
echo "--- burst pair=$m" >> $O;  B200RL_FUSED_PAIR=$m timeout 120 python scripts/time_fused.py 40 >> $O 2>&1
done
cat $O
