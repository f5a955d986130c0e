"""Short single-GPU driver for ncu: one forward (entropy + min/max) and one backward of the
log-prob kernels on T x 151936 bf16 logits, plus GAE and interval copies at bench sizes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import areal_vllm_b200 as pkg

T = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
what = sys.argv[2] if len(sys.argv) > 2 else "h1"
dev = torch.device("cuda:0")
V = 151936
if what == "h1":
    gen = torch.Generator(device=dev).manual_seed(0)
    x = torch.empty(T, V, dtype=torch.bfloat16, device=dev)
    for lo in range(0, T, 2048):
        x[lo:lo + 2048] = (torch.randn(min(2048, T - lo), V, generator=gen, device=dev) * 2).bfloat16()
    y = torch.randint(0, V, (T,), generator=gen, device=dev)
    g = torch.randn(T, generator=gen, device=dev)
    for _ in range(2):
        logp, ent, lse, mx, mn = pkg.ops.logprob_entropy_fwd(x, y, 1.0, True, True)
        logp2, _, lse2, _, _ = pkg.ops.logprob_entropy_fwd(x, y, 1.0, False, False)
        d = pkg.ops.logprob_entropy_bwd(x, y, lse, ent, g, None, 1.0)
        d2 = pkg.ops.logprob_entropy_bwd(x, y, lse, ent, g, 0.01 * g, 1.0)
    torch.cuda.synchronize()
    print("ok", float(logp.sum()), float(d.float().abs().sum()))
elif what == "fused":
    from areal_vllm_b200.actor import grpo_loss_fn_fused
    gen = torch.Generator(device=dev).manual_seed(0)
    x = torch.empty(T, V, dtype=torch.bfloat16, device=dev)
    for lo in range(0, T, 2048):
        x[lo:lo + 2048] = (torch.randn(min(2048, T - lo), V, generator=gen, device=dev) * 2).bfloat16()
    y = torch.randint(0, V, (T,), generator=gen, device=dev)
    data = dict(input_ids=y, logprobs=-torch.rand(T, device=dev) * 4, advantages=torch.randn(T, device=dev),
                loss_mask=(torch.rand(T, device=dev) < 0.7).float(), prox_logp=-torch.rand(T, device=dev) * 4)
    for _ in range(3):
        xg = x.detach().requires_grad_(True)
        loss = grpo_loss_fn_fused(xg, data, 1.0, 0.4, None, None, 5.0, loss_scale=torch.tensor(1.0, device=dev))
        (d,) = torch.autograd.grad(loss, xg)
    torch.cuda.synchronize()
    print("ok", float(loss), float(d.float().abs().sum()))
elif what == "gae":
    rng = np.random.default_rng(0)
    B = 65536
    lens = rng.integers(64, 2049, B)
    total = int(lens.sum())
    r = torch.randn(total, device=dev); v = torch.randn(total + B, device=dev)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
    b = torch.ones(B, dtype=torch.bool, device=dev)
    for _ in range(2):
        a1 = pkg.ppo_functional.gae_1d_nolp_misalign(r, v, cu, b, 0.99, 0.95, mode="scan")
        a2 = pkg.ppo_functional.gae_1d_nolp_misalign(r, v, cu, b, 1.0, 1.0, mode="serial")
    torch.cuda.synchronize()
    print("ok", total)
elif what == "gae_team":
    # few long sequences: the team kernel (the longest chain is the run time)
    rng = np.random.default_rng(120)
    B = 120
    lens = rng.integers(1024, 16385, B)
    total = int(lens.sum())
    r = torch.randn(total, device=dev); v = torch.randn(total + B, device=dev)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
    b = torch.ones(B, dtype=torch.bool, device=dev)
    for _ in range(3):
        a2 = pkg.ppo_functional.gae_1d_nolp_misalign(r, v, cu, b, 1.0, 1.0, mode="serial_team")
    torch.cuda.synchronize()
    print("ok", total)
