"""GAE kernels by mode over a sweep of batch sizes (length law 64..2048 and a short-sequence law)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from areal_vllm_b200 import _lib
dev = torch.device("cuda:0")
flush = torch.empty(256 * 2**20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=8, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))
MODES = {"scan": 2, "warp": 3, "relay": 7, "slots": 6, "team": 8}
def case(B, lo, hi, gl=(1.0, 1.0)):
    rng = np.random.default_rng(B)
    lens = rng.integers(lo, hi + 1, B); total = int(lens.sum())
    r = torch.randn(total, device=dev); v = torch.randn(total + B, device=dev)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
    b = torch.ones(B, dtype=torch.bool, device=dev).view(torch.uint8)
    adv = torch.empty_like(r); ret = torch.empty_like(r)
    ws = _lib.workspace(_lib.lib.b200rl_gae_workspace_bytes(B, total), dev)
    out = []
    for name, m in MODES.items():
        def run():
            rc = _lib.lib.b200rl_gae_1d_nolp_misalign(_lib.ptr(r), _lib.ptr(v), _lib.ptr(cu), _lib.ptr(b), _lib.ptr(adv),
                                                      _lib.ptr(ret), B, total, gl[0], gl[1], m, _lib.ptr(ws), ws.numel(),
                                                      _lib.stream_handle(dev))
            assert rc == 0, _lib.last_error()
        ms = timeit(run)
        out.append(f"{name} {ms*1e3:8.1f} us ({16.0*total/ms/1e6/6529.4:5.3f})")
    print(f"B={B:7d} len[{lo},{hi}] tok={total:10d}  " + "  ".join(out))
if "--small" not in sys.argv:
    for B in (256, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072):
        case(B, 64, 2048)
    for B in (8192, 65536, 500000):
        case(B, 1, 64)
    case(4096, 1024, 16384)
# few sequences: the longest chain is the run time (team vs warp)
case(120, 1024, 16384)
case(512, 1024, 16384)
case(1000, 1024, 16384)
case(16, 64, 2048)
case(512, 64, 2048)
case(1000, 512, 1024)
