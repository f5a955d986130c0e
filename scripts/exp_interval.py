import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import areal_vllm_b200 as pkg
dev = torch.device("cuda:0")
flush = torch.empty(256 * 2**20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=6, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))
size = 7_615_616_512 // 2
src = torch.empty(size, dtype=torch.bfloat16, device=dev); src.view(torch.int16).random_()
def case(name, iv):
    ivc = torch.from_numpy(iv).to(dev); total = int((iv[:,1]-iv[:,0]).sum())
    pkg.flatten_param.clear_plan_cache()
    out = pkg.flatten_param.slice_intervals(src, intervals_cuda=ivc, output_size=total)
    ms = timeit(lambda: pkg.flatten_param.slice_intervals(src, intervals_cuda=ivc, output_size=total))
    dst = torch.empty_like(src)
    ms2 = timeit(lambda: pkg.flatten_param.set_intervals(out, dst, intervals_cuda=ivc))
    gb = 4.0 * total / 1e9
    n_d = max(1, len(iv) // 3); per = total // n_d
    d_iv = np.array([(i * (size // n_d), i * (size // n_d) + (per if i < n_d - 1 else total - per * (n_d - 1))) for i in range(n_d)], dtype=np.int64)
    dvc = torch.from_numpy(d_iv).to(dev)
    ms3 = timeit(lambda: pkg.flatten_param.move_intervals(src, dst, ivc, dvc, total))
    print(f"{name:50s} slice {gb/ms*1e3:7.0f} GB/s  set {gb/ms2*1e3:7.0f} GB/s  move {gb/ms3*1e3:7.0f} GB/s")
    del out, dst
n = 50000
s = size // n
base = np.array([(i * s, i * s + s // 2) for i in range(n)], dtype=np.int64)
case("cfg5/2 sequential order (misaligned)", base)
sh = base.copy(); np.random.default_rng(5).shuffle(sh)
case("cfg5/2 shuffled (misaligned)", sh)
s8 = (s // 16) * 16
al = np.array([(i * s8, i * s8 + s8 // 2) for i in range(n)], dtype=np.int64)
case("aligned sequential", al)
sh = al.copy(); np.random.default_rng(5).shuffle(sh)
case("aligned shuffled", sh)
one = np.array([(0, size // 2)], dtype=np.int64)
case("one interval (plain copy)", one)
d = torch.empty(size // 2, dtype=torch.bfloat16, device=dev)
ms = timeit(lambda: d.copy_(src[: size // 2]))
print(f"torch copy_ same bytes: {4.0*(size//2)/1e9/ms*1e3:7.0f} GB/s")
