#!/bin/bash
# round 2, N GPUs: bench.py under torchrun (parity block included) + the two *_dist tests over NCCL
N=${1:-2}
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 \
  bench.py --gpus $N --steps ${2:-5} --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/bench_n$N.json; grep -v "^W1107\|^\*\*\*\|^$" gpurun_out/bench_n$N.err | tail -15
timeout 600 python -m pytest tests/test_move_intervals_dist.py tests/test_vocab_parallel_dist.py -q -m gpu -p no:cacheprovider 2>&1 | tail -3
