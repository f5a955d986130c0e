import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import areal_vllm_b200 as pkg
from areal_vllm_b200.actor import grpo_loss_fn_fused
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
data = dict(input_ids=y, logprobs=-torch.rand(T, device=dev) * 4, advantages=torch.randn(T, device=dev),
            loss_mask=(torch.rand(T, device=dev) < 0.7).float(), prox_logp=-torch.rand(T, device=dev) * 4)
ls = torch.tensor(1.0, device=dev)
def fused(x):
    xg = x.detach().requires_grad_(True)
    loss = grpo_loss_fn_fused(xg, data, 1.0, 0.4, None, None, 5.0, loss_scale=ls)
    torch.autograd.grad(loss, xg)
for i in range(3): fused(xs[i % 3])
torch.cuda.synchronize()
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 30
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(iters): fused(xs[i % 3])
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(f"fused step T={T}: {ms*1e3:8.1f} us  {T/ms/1e3:6.3f} M tok/s  HBM(4V) {4.0*T*V/ms/1e6:6.0f} GB/s  6V-equiv {6.0*T*V/ms/1e6:6.0f} GB/s")
