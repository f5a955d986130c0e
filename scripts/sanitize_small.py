"""Small invocation of every kernel (ragged shapes included) for compute-sanitizer runs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import areal_vllm_b200 as pkg
from areal_vllm_b200 import functional as BF, ppo_functional as PF, flatten_param as FP, actor as AC
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
for dtype, V in ((torch.bfloat16, 4104), (torch.bfloat16, 151936), (torch.float32, 1028), (torch.float16, 2048), (torch.bfloat16, 1001)):
    T = 9 if V > 100000 else 37
    x = (torch.randn(T, V, generator=g) * 2).to(dtype).to(dev).requires_grad_(True)
    y = torch.randint(0, V, (T,), generator=g).to(dev)
    lp, ent, mx, mn = BF.gather_logprobs_entropy_minmax(x, y, 0.9)
    data = dict(input_ids=y, logprobs=lp.detach() - 0.1, advantages=torch.randn(T, device=dev),
                loss_mask=(torch.rand(T, device=dev) < 0.7).float(), prox_logp=lp.detach() + 0.05)
    loss, st = BF.ppo_actor_loss_fn(lp, data["prox_logp"], data["logprobs"], data["advantages"], 0.2, data["loss_mask"].bool(), c_clip=3.0, behav_imp_weight_cap=5.0)
    (loss + 0.01 * ent.mean()).backward()
    x2 = x.detach().clone().requires_grad_(True)
    l2 = AC.grpo_loss_fn_fused(x2, data, 0.9, 0.2, 0.28, 3.0, 5.0, loss_scale=torch.tensor(0.5, device=dev))
    (l2 * 2.0).backward()
rng = np.random.default_rng(0)
lens = np.array([0, 1, 1024, 0, 2048, 1, 1023, 3000, 5, 0], dtype=np.int64)
total, B = int(lens.sum()), len(lens)
r = torch.randn(total, device=dev); v = torch.randn(total + B, device=dev)
cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
b = torch.from_numpy(rng.random(B) < 0.5).to(dev)
for mode in ("serial", "scan"):
    PF.gae_1d_nolp_misalign(r, v, cu, b, 0.99, 0.95, mode=mode)
AC.padded_gae(torch.randn(5, 33, device=dev), torch.randn(5, 33, device=dev), torch.ones(5, dtype=torch.bool, device=dev), 1.0, 1.0)
class Cfg: mean_level, std_level, group_size = "group", "batch", 4
AC.AdvNorm(Cfg)(torch.randn(18, 12, device=dev), torch.ones(18, 12, device=dev))
BF.masked_normalization(torch.randn(6, 10, device=dev), torch.ones(6, 10, device=dev), dim=(0,))
size = 70001
iv = []
pos = 0
for k in range(400):
    pos += int(rng.integers(0, 23)); ln = int(rng.integers(0, 300)); iv.append((pos, pos + ln)); pos += ln
iv.append((pos + 3, pos + 3 + 30001)); iv = np.array(iv, dtype=np.int64)[rng.permutation(len(iv))]
for dt in (torch.bfloat16, torch.float32):
    src = torch.randn(size, device=dev).to(dt)
    out = FP.slice_intervals(src, intervals_cuda=torch.from_numpy(iv).to(dev), output_size=int((iv[:, 1] - iv[:, 0]).sum()))
    FP.set_intervals(out, torch.zeros_like(src), intervals_cuda=torch.from_numpy(iv).to(dev))
torch.cuda.synchronize()
print("sanitize_small ok")
