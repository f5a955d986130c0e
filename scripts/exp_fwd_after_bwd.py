"""Is the forward slower right after a backward (DRAM write -> read turnaround, dirty L2 tail) or is
the in-step gap host/launch overhead? Times the raw forward launcher in three contexts."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from areal_vllm_b200 import ops
dev = torch.device("cuda:0")
T, V = 10240, 151936
bufs = [(torch.randn(T, V, device=dev) * 2).bfloat16() for _ in range(4)]
y = torch.randint(0, V, (T,), device=dev)
g = torch.randn(T, device=dev)
def ev(): return torch.cuda.Event(enable_timing=True)
def run(pattern, iters=12):
    ts = []
    for it in range(iters + 3):
        x = bufs[it % 4]
        if pattern == "bwd_then_fwd":
            o = ops._raw_logprob_entropy_fwd(bufs[(it + 1) % 4], y, 1.0, True, True)
            d = ops._raw_logprob_entropy_bwd(bufs[(it + 1) % 4], y, o[2], o[1], g, None, 1.0)
        e0, e1 = ev(), ev()
        e0.record()
        out = ops._raw_logprob_entropy_fwd(x, y, 1.0, True, True)
        e1.record()
        if pattern == "fwd_then_small":
            for _ in range(4):
                _ = out[0] * 2.0
        torch.cuda.synchronize()
        if it >= 3: ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    print(f"{pattern:16s} fwd {ms*1e3:7.1f} us  {2.0*T*V/ms/1e6:7.0f} GB/s")
run("fwd_only"); run("bwd_then_fwd"); run("fwd_only"); run("bwd_then_fwd")
