#!/bin/bash
N=${1:-2}
for T in 2048 8192; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 scripts/tp_check.py $T 2>&1 | grep '^{' | python -c "
import sys, json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('tp', d['tp'], 'T', d['T'], 'ok', d['ok'], '| all_gather %.3f ms %.0f GB/s | peer_stores %s' % (d['all_gather']['fwd_bwd_ms'], d['all_gather']['shard_GBps'], ('%.3f ms %.0f GB/s ok=%s' % (d['peer_stores']['fwd_bwd_ms'], d['peer_stores']['shard_GBps'], d['peer_stores']['ok'])) if 'fwd_bwd_ms' in d['peer_stores'] else d['peer_stores']))
"
done
