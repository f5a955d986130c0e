import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from areal_vllm_b200 import _lib
dev = torch.device("cuda:0")
for n in (65536, 1048576):
    r = torch.randn(n, device=dev); v = torch.randn(n + 1, device=dev)
    cu = torch.tensor([0, n], dtype=torch.int32, device=dev); b = torch.ones(1, dtype=torch.uint8, device=dev)
    adv = torch.empty_like(r); ret = torch.empty_like(r)
    ws = _lib.workspace(_lib.lib.b200rl_gae_workspace_bytes(1, n), dev)
    for name, m in (("warp", 3), ("relay", 7), ("team", 8)):
        def run():
            rc = _lib.lib.b200rl_gae_1d_nolp_misalign(_lib.ptr(r), _lib.ptr(v), _lib.ptr(cu), _lib.ptr(b), _lib.ptr(adv), _lib.ptr(ret), 1, n, 1.0, 1.0, m, _lib.ptr(ws), ws.numel(), _lib.stream_handle(dev))
            assert rc == 0
        run(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"n={n} {name}: {ms*1e3:.1f} us  {ms*1e6/n:.2f} ns/token  ~{ms*1e6/n*1.92:.2f} cycles/token @1.92GHz")
