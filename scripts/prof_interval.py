"""cfg5-shaped interval slice / set / move, twice each (for ncu: scripts/gpu_prof_interval.sh)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import areal_vllm_b200 as pkg  # noqa: E402

dev = torch.device("cuda:0")
size, n = 7_615_616_512 // 2, 100000 // 2
s_ = size // n
iv = np.array([(i * s_, i * s_ + s_ // 2) for i in range(n)], dtype=np.int64)
np.random.default_rng(5).shuffle(iv)
src = torch.empty(size, dtype=torch.bfloat16, device=dev)
src.view(torch.int16).random_()
dst = torch.empty(size, dtype=torch.bfloat16, device=dev)
ivc = torch.from_numpy(iv).to(dev)
total = int((iv[:, 1] - iv[:, 0]).sum())
n_d = n // 3
per = total // n_d
d_iv = np.array([(i * (size // n_d), i * (size // n_d) + (per if i < n_d - 1 else total - per * (n_d - 1)))
                 for i in range(n_d)], dtype=np.int64)
dvc = torch.from_numpy(d_iv).to(dev)
for _ in range(2):
    packed = pkg.flatten_param.slice_intervals(src, intervals_cuda=ivc, output_size=total)
    pkg.flatten_param.set_intervals(packed, dst, intervals_cuda=ivc)
    pkg.flatten_param.move_intervals(src, dst, ivc, dvc, total)
torch.cuda.synchronize()
print("ok", total)
