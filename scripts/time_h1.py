"""CUDA-event timing of the individual H1 kernels (device-resident inputs, > L2)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import areal_vllm_b200 as pkg

dev = torch.device("cuda:0")
T, V = 10240, 151936
gen = torch.Generator(device=dev).manual_seed(0)
xs = []
for b in range(3):
    x = torch.empty(T, V, dtype=torch.bfloat16, device=dev)
    for lo in range(0, T, 2048):
        x[lo:lo + 2048] = (torch.randn(2048, V, generator=gen, device=dev) * 2).bfloat16()
    xs.append(x)
y = torch.randint(0, V, (T,), generator=gen, device=dev)
g = torch.randn(T, generator=gen, device=dev)
PEAK = 6529.4

def bench(name, fn, bytes_per_call, iters=12):
    for i in range(3):
        fn(xs[i % 3])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(xs[i % 3])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    gbs = bytes_per_call / ms / 1e6
    print(f"{name:28s} {ms*1e3:9.1f} us  {gbs:8.0f} GB/s  {gbs/PEAK:6.3f} of measured copy peak")

lse = ent = None
def fwd(ent_, mm):
    def f(x):
        global lse, ent
        out = pkg.ops.logprob_entropy_fwd(x, y, 1.0, ent_, mm)
        lse, ent = out[2], out[1]
    return f
bench("fwd (logp only)", fwd(False, False), 2.0 * T * V)
bench("fwd (+entropy)", fwd(True, False), 2.0 * T * V)
bench("fwd (+entropy,+minmax)", fwd(True, True), 2.0 * T * V)
bench("bwd (g_ent = None)", lambda x: pkg.ops.logprob_entropy_bwd(x, y, lse, ent, g, None, 1.0), 4.0 * T * V)
bench("bwd (with g_ent)", lambda x: pkg.ops.logprob_entropy_bwd(x, y, lse, ent, g, 0.01 * g, 1.0), 4.0 * T * V)
d = torch.empty_like(xs[0])
bench("torch copy_ (reference)", lambda x: d.copy_(x), 4.0 * T * V)
bench("torch sum fp32 (read-only)", lambda x: x.view(torch.int32).sum(), 2.0 * T * V)

# fp32 logits (Megatron path): 4*V bytes per token forward, 8*V backward
T32 = 4096
x32 = [xs[i][:T32].float() for i in range(2)]
y32, g32 = y[:T32], g[:T32]
def bench32(name, fn, bytes_per_call, iters=8):
    for i in range(2): fn(x32[i % 2])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters): fn(x32[i % 2])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"{name:28s} {ms*1e3:9.1f} us  {bytes_per_call/ms/1e6:8.0f} GB/s  {bytes_per_call/ms/1e6/PEAK:6.3f} of measured copy peak")
def f32(x):
    global lse, ent
    out = pkg.ops.logprob_entropy_fwd(x, y32, 1.0, True, False); lse, ent = out[2], out[1]
bench32("fp32 fwd (+entropy)", f32, 4.0 * T32 * V)
bench32("fp32 bwd", lambda x: pkg.ops.logprob_entropy_bwd(x, y32, lse, ent, g32, None, 1.0), 8.0 * T32 * V)

# fused GRPO step: 4*V bytes of HBM traffic per token (bf16)
from areal_vllm_b200.actor import grpo_loss_fn_fused, grpo_loss_fn
Tm = 10240
data = dict(input_ids=y[:Tm], logprobs=-torch.rand(Tm, device=dev) * 4, advantages=torch.randn(Tm, device=dev),
            loss_mask=(torch.rand(Tm, device=dev) < 0.7).float(), prox_logp=-torch.rand(Tm, device=dev) * 4)
ls = torch.tensor(0.5, device=dev)
def fused(x):
    xg = x[:Tm].detach().requires_grad_(True)
    loss = grpo_loss_fn_fused(xg, data, 1.0, 0.4, None, None, 5.0, loss_scale=ls)
    torch.autograd.grad(loss, xg)
def separate(x):
    xg = x[:Tm].detach().requires_grad_(True)
    loss = grpo_loss_fn(xg, data, 1.0, 0.4, None, None, 5.0)
    torch.autograd.grad(loss * ls, xg)
bench("separate fwd+loss+bwd (6V)", separate, 6.0 * Tm * V)
bench("FUSED step (bytes=4V)", fused, 4.0 * Tm * V)
bench("FUSED step (as 6V-equiv)", fused, 6.0 * Tm * V)
