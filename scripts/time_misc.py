"""CUDA-event timing of GAE, interval copy, PPO loss and AdvNorm kernels at BASELINE sizes."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import areal_vllm_b200 as pkg
from areal_vllm_b200 import _lib

dev = torch.device("cuda:0")
PEAK = 6529.4
flush = torch.empty(256 * 2**20, dtype=torch.uint8, device=dev)

def timeit(fn, iters=20, warm=3, flush_l2=True):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush_l2:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))

def gae_case(name, B, lo, hi, seed):
    rng = np.random.default_rng(seed)
    lens = rng.integers(lo, hi + 1, B)
    total = int(lens.sum())
    r = torch.randn(total, device=dev); v = torch.randn(total + B, device=dev)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
    b = torch.ones(B, dtype=torch.bool, device=dev)
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    ws = _lib.workspace(_lib.lib.b200rl_gae_workspace_bytes(B, total), dev)
    st = _lib.stream_handle(dev)
    for mode, mname, g, l in ((2, "scan", 0.99, 0.95), (1, "serial", 1.0, 1.0)):
        def f():
            rc = _lib.lib.b200rl_gae_1d_nolp_misalign(_lib.ptr(r), _lib.ptr(v), _lib.ptr(cu), _lib.ptr(b), _lib.ptr(adv),
                                                      _lib.ptr(ret), B, total, g, l, mode, _lib.ptr(ws), ws.numel(), st)
            assert rc == 0
        ms = timeit(f)
        gbs = 16.0 * total / ms / 1e6
        print(f"GAE {name:22s} {mname:6s} B={B:6d} tokens={total:9d}  {ms*1e3:9.1f} us  {total/ms/1e6:9.2f} Gtok/s  {gbs:7.0f} GB/s ({gbs/PEAK:5.3f})")

from oracle import ref_loader as R
REF_GAE, REF_IV = R.ref_cugae_cuda(), R.ref_interval_op_cuda()

def ref_gae_case(name, B, lo, hi, seed):
    if REF_GAE is None:
        return
    rng = np.random.default_rng(seed)
    lens = rng.integers(lo, hi + 1, B)
    total = int(lens.sum())
    r = torch.randn(total, device=dev); v = torch.randn(total + B, device=dev)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
    b = torch.ones(B, dtype=torch.bool, device=dev)
    ms = timeit(lambda: REF_GAE.gae_1d_nolp_misalign(r, v, cu, b, 0.99, 0.95), iters=8)
    print(f"GAE {name:22s} REFERENCE gae.cu (sm_100a build)      {ms*1e3:9.1f} us  {total/ms/1e6:9.2f} Gtok/s")

for args in (("cfg1 256x[64,2048]", 256, 64, 2048, 1), ("cfg4 120x[1k,16k]", 120, 1024, 16384, 4),
             ("large 65536x[64,2048]", 65536, 64, 2048, 2)):
    ref_gae_case(*args)
gae_case("cfg1 256x[64,2048]", 256, 64, 2048, 1)
gae_case("cfg4 120x[1k,16k]", 120, 1024, 16384, 4)
gae_case("large 65536x[64,2048]", 65536, 64, 2048, 2)
gae_case("short 500k x [1,64]", 500000, 1, 64, 3)

# ---- intervals ---------------------------------------------------------------------------
def interval_case(name, size, iv, dtype=torch.bfloat16):
    es = torch.empty((), dtype=dtype).element_size()
    src = torch.empty(size, dtype=dtype, device=dev)
    src.view(torch.int16 if es == 2 else torch.int32).random_()
    ivc = torch.from_numpy(iv).to(dev)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    pkg.flatten_param.clear_plan_cache()
    t0 = time.perf_counter()
    out = pkg.flatten_param.slice_intervals(src, intervals_cuda=ivc, output_size=total)
    torch.cuda.synchronize()
    ms_s = timeit(lambda: pkg.flatten_param.slice_intervals(src, intervals_cuda=ivc, output_size=total), iters=8)
    dst = torch.empty_like(src)
    ms_t = timeit(lambda: pkg.flatten_param.set_intervals(out, dst, intervals_cuda=ivc), iters=8)
    def replan():
        pkg.flatten_param.clear_plan_cache()
        pkg.flatten_param.slice_intervals(src, intervals_cuda=ivc, output_size=total)
    ms_p = timeit(replan, iters=5)
    gb = 2.0 * total * es / 1e9
    if REF_IV is not None and dtype == torch.bfloat16:
        mx = int((iv[:, 1] - iv[:, 0]).max())
        ms_r = timeit(lambda: REF_IV.slice_intervals_bf16(src, ivc, total, mx), iters=3, warm=1)
        ms_rs = timeit(lambda: REF_IV.set_intervals_bf16(out, dst, ivc, mx), iters=3, warm=1)
        print(f"INTERVAL {name:34s} REFERENCE interval_op.cu: slice {ms_r*1e3:9.1f} us {gb/ms_r*1e3:7.0f} GB/s  set {ms_rs*1e3:9.1f} us {gb/ms_rs*1e3:7.0f} GB/s")
    print(f"INTERVAL {name:34s} N={len(iv):7d} copied={total*es/1e9:7.3f} GB  slice {ms_s*1e3:9.1f} us {gb/ms_s*1e3:7.0f} GB/s ({gb/ms_s*1e3/PEAK:5.3f})"
          f"  set {ms_t*1e3:9.1f} us {gb/ms_t*1e3:7.0f} GB/s ({gb/ms_t*1e3/PEAK:5.3f})  slice+plan {ms_p*1e3:9.1f} us")
    del src, dst, out

size = 7_615_616_512
n = 100000
s = size // n
iv = np.array([(i * s, i * s + s // 2) for i in range(n)], dtype=np.int64)
np.random.default_rng(5).shuffle(iv)
interval_case("cfg5 make_intervals(7.6e9, 1e5)", size, iv)
# realistic TP reshard mix: per layer 3584 rows of 1792 and 3584 rows of 9472 (row-linear splits) + long merged runs
ivs, pos = [], 0
for layer in range(28):
    for rows, ln, stride in ((3584, 1792, 3584), (3584, 9472, 18944)):
        st_ = pos + np.arange(rows, dtype=np.int64) * stride
        ivs.append(np.stack([st_, st_ + ln], 1)); pos += rows * stride
    for ln in (12_845_056, 67_895_296):
        ivs.append(np.array([[pos, pos + ln]], dtype=np.int64)); pos += ln
iv2 = np.concatenate(ivs)
interval_case("TP1->2 reshard mix (Qwen2.5-7B)", int(pos), iv2)
rng = np.random.default_rng(9)
st_ = np.sort(rng.choice(400_000_000, 200000, replace=False)).astype(np.int64) * 2 + 1
ln = rng.integers(1, 1500, 200000)
ln = np.minimum(ln, np.diff(np.append(st_, 800_000_001)) - 1)
iv3 = np.stack([st_, st_ + ln], 1)
interval_case("adversarial odd offsets, len<1500", 800_000_100, iv3)

# ---- PPO loss + advnorm -------------------------------------------------------------------
for T in (10240, 32768, 1 << 20):
    g = torch.Generator(device=dev).manual_seed(0)
    lp = -torch.rand(T, device=dev, generator=g); prox = lp + 0.02 * torch.randn(T, device=dev, generator=g)
    old = prox + 0.05 * torch.randn(T, device=dev, generator=g); adv = torch.randn(T, device=dev, generator=g)
    mask = torch.rand(T, device=dev, generator=g) < 0.7
    lpg = lp.clone().requires_grad_(True)
    def fb():
        loss, stat = pkg.functional.ppo_actor_loss_fn(lpg, prox, old, adv, 0.2, mask, eps_clip_higher=0.28, c_clip=3.0, behav_imp_weight_cap=5.0)
        torch.autograd.grad(loss, lpg)
    ms = timeit(fb, flush_l2=False)
    print(f"PPO loss fwd+bwd via public API T={T:8d}: {ms*1e3:8.1f} us")
for T, V in ((10240, 2048), (2048, 2048)):
    g = torch.Generator(device=dev).manual_seed(1)
    xl = (torch.randn(T, V, device=dev, generator=g) * 2).bfloat16().requires_grad_(True)
    data = dict(input_ids=torch.randint(0, V, (T,), device=dev, generator=g), logprobs=-torch.rand(T, device=dev, generator=g),
                advantages=torch.randn(T, device=dev, generator=g), loss_mask=(torch.rand(T, device=dev, generator=g) < 0.7).float(),
                prox_logp=-torch.rand(T, device=dev, generator=g))
    def gl(stats):
        loss = pkg.actor.grpo_loss_fn(xl, data, 1.0, 0.2, 0.28, 3.0, 5.0, stats=stats)
        torch.autograd.grad(loss, xl)
    print(f"grpo_loss_fn fwd+bwd via public API T={T:6d} V={V} (small V: host-bound), stats dict: {timeit(lambda: gl({}), flush_l2=False)*1e3:8.1f} us"
          f"   no stats: {timeit(lambda: gl(None), flush_l2=False)*1e3:8.1f} us")
class Cfg: mean_level, std_level, group_size = "group", "group", 8
x = torch.randn(512, 1024, device=dev); m = (torch.rand(512, 1024, device=dev) < 0.6).float()
an = pkg.actor.AdvNorm(Cfg)
print(f"AdvNorm group/group [512,1024]: {timeit(lambda: an(x, m), flush_l2=False)*1e3:8.1f} us")
Cfg.mean_level = Cfg.std_level = "batch"
an = pkg.actor.AdvNorm(Cfg)
print(f"AdvNorm batch/batch [512,1024]: {timeit(lambda: an(x, m), flush_l2=False)*1e3:8.1f} us")
