"""Checksums of every output of the fused GRPO step on a few shapes; run once with
B200RL_FUSED_PAIR=0 and once with =1 and compare the printed lines (they must be identical:
the two forms do the same arithmetic in the same order)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from areal_vllm_b200.actor import grpo_loss_fn_fused

dev = torch.device("cuda:0")


def digest(t):
    if t.dtype == torch.bool:
        t = t.to(torch.uint8)
    b = t.contiguous().view(torch.uint8).reshape(-1).to(torch.int64)
    w = torch.arange(b.numel(), device=b.device, dtype=torch.int64) % 65521 + 1
    return [int(b.sum()), int((b * w).sum() % (1 << 61))]


def one(T, V, dtype, seed):
    gen = torch.Generator(device=dev).manual_seed(seed)
    x = (torch.randn(T, V, generator=gen, device=dev) * 2).to(dtype)
    ids = torch.randint(0, V, (T,), generator=gen, device=dev)
    data = dict(input_ids=ids, logprobs=-torch.rand(T, generator=gen, device=dev) * 4,
                advantages=torch.randn(T, generator=gen, device=dev),
                loss_mask=(torch.rand(T, generator=gen, device=dev) < 0.7).float(),
                prox_logp=-torch.rand(T, generator=gen, device=dev) * 4)
    xg = x.detach().requires_grad_(True)
    stats = {}
    loss = grpo_loss_fn_fused(xg, data, 0.9, 0.2, 0.28, 3.0, 5.0, loss_scale=torch.tensor(0.5, device=dev),
                              stats=stats)
    (g,) = torch.autograd.grad(loss, xg)
    torch.cuda.synchronize()
    out = {"T": T, "V": V, "dtype": str(dtype), "loss": loss.item().hex(), "dlogits": digest(g)}
    for k in sorted(stats):  # log-probs, entropy, row min/max, every logged tensor of the loss
        if isinstance(stats[k], torch.Tensor):
            out[k] = digest(stats[k])
    return out


if __name__ == "__main__":
    shapes = [(1, 4096, torch.bfloat16), (2, 4096, torch.bfloat16), (3, 1024, torch.bfloat16),
              (297, 151936, torch.bfloat16), (1000, 32000, torch.float32), (5001, 1024, torch.float16),
              (149, 152064, torch.bfloat16)]
    for i, (T, V, dt) in enumerate(shapes):
        print(json.dumps(one(T, V, dt, i)))
