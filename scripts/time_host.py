"""Host-side cost per call of the drop-in functions (CPU time, tiny T so the GPU is never the limit)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import areal_vllm_b200 as pkg
from areal_vllm_b200 import ops, functional as BF
dev = torch.device("cuda:0")
T, V = 256, 4096
x = torch.randn(T, V, device=dev).bfloat16(); y = torch.randint(0, V, (T,), device=dev)
prox = -torch.rand(T, device=dev); old = prox.clone(); adv = torch.randn(T, device=dev); mask = torch.rand(T, device=dev) < 0.7
def t(name, fn, n=300):
    for _ in range(20): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    t1 = time.perf_counter(); torch.cuda.synchronize()
    print(f"{name:46s} {1e6*(t1-t0)/n:8.1f} us/call (host)")
t("raw logprob fwd", lambda: ops._raw_logprob_entropy_fwd(x, y, 1.0, True, True))
t("custom-op logprob fwd", lambda: ops.logprob_entropy_fwd_op(x, y, 1.0, True, True))
lp = ops.logprob_entropy_fwd(x, y, 1.0, False, False)[0]
t("raw ppo_loss_fwd", lambda: ops._raw_ppo_loss_fwd(lp, prox, old, adv, mask, 0.8, 1.2, -1.0, 5.0, None, None, None))
t("custom-op ppo_loss_fwd", lambda: ops.ppo_loss_fwd_op(lp, prox, old, adv, mask, 0.8, 1.2, -1.0, 5.0, None, None, None))
def full():
    xg = x.detach().requires_grad_(True)
    l, e = BF.gather_logprobs_entropy(xg, y, 1.0)
    loss, st = BF.ppo_actor_loss_fn(l, prox, old, adv, 0.2, mask, behav_imp_weight_cap=5.0)
    torch.autograd.grad(loss, xg)
t("drop-in fwd+loss+bwd (eager fast path)", full)
data = dict(input_ids=y, logprobs=old, advantages=adv, loss_mask=mask.float(), prox_logp=prox)
def full2():
    xg = x.detach().requires_grad_(True)
    loss = pkg.actor.grpo_loss_fn(xg, data, 1.0, 0.4, None, None, 5.0)
    torch.autograd.grad(loss, xg)
t("grpo_loss_fn + backward", full2)
class Cfg: mean_level, std_level, group_size = "group", "group", 8
an = pkg.actor.AdvNorm(Cfg); a = torch.randn(64, 128, device=dev); m = torch.ones(64, 128, device=dev)
t("AdvNorm group/group", lambda: an(a, m))
# newer entry points
from areal_vllm_b200 import mcore_functional as MF, ppo_functional as PF, lm_engine as LE
def tp1():
    xg = x.detach().requires_grad_(True)
    l, e = MF.vocab_parallel_logprobs_entropy(xg, y, 1.0)
    torch.autograd.grad(l.sum(), xg)
t("vocab_parallel fwd+bwd (one shard, no group)", tp1)
val = torch.randn(T, device=dev)
def crit():
    vg = val.detach().requires_grad_(True)
    l, s = PF.critic_loss_fn(vg, val + 0.1, val - 1.0, 0.2, mask)
    torch.autograd.grad(l, vg)
t("critic_loss_fn + backward", crit)
cu = torch.tensor([0, 100, 256], dtype=torch.int32, device=dev)
sc = torch.randn(2, device=dev); ne = torch.zeros(2, dtype=torch.bool, device=dev)
t("get_packed_rewards", lambda: PF.get_packed_rewards(0.1, 5.0, prox, old, sc, cu, ne))
inp = dict(input_ids=y, cu_seqlens=cu, loss_mask=mask.int())
def sft():
    xg = x.detach().requires_grad_(True)
    torch.autograd.grad(LE.compute_packed_sft_loss(xg, inp), xg)
t("compute_packed_sft_loss + backward", sft)
def sftf():
    xg = x.detach().requires_grad_(True)
    torch.autograd.grad(LE.compute_packed_sft_loss_fused(xg, inp), xg)
t("compute_packed_sft_loss_fused + backward", sftf)
