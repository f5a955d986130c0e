"""Vocabulary-parallel log-prob over a real NCCL group: run under torchrun with N GPUs, e.g.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29533 scripts/tp_check.py
Every rank builds the same full logits, keeps its column shard, and checks the sharded result and
gradient against the whole-vocabulary kernel run locally. Also times the sharded forward+backward.
Prints one JSON line on rank 0."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_vllm_b200 import functional, mcore_functional  # noqa: E402


def _one(dev, rank, world, T, V, group, full, labels, g1, g2):
    Vs = V // world
    shard = full[:, rank * Vs:(rank + 1) * Vs].contiguous().requires_grad_(True)
    lp, ent = mcore_functional.vocab_parallel_logprobs_entropy(shard, labels, 0.9, group=group)
    ((lp * g1).sum() + (ent * g2).sum()).backward()
    ref_in = full.clone().requires_grad_(True)
    lp_r, ent_r = functional.gather_logprobs_entropy(ref_in, labels, 0.9)
    ((lp_r * g1).sum() + (ent_r * g2).sum()).backward()
    err = dict(
        logp=float((lp.detach() - lp_r.detach()).abs().max()),
        ent=float((ent.detach() - ent_r.detach()).abs().max()),
        grad=float((shard.grad.float() - ref_in.grad[:, rank * Vs:(rank + 1) * Vs].float()).abs().max()),
    )
    # bars: logp / entropy 1e-3 relative (north_star); dlogits within one bf16 ulp of the
    # whole-vocabulary kernel's gradient (2^-7 relative to the element, like tests/test_logprob_gpu.py)
    gs, gr = shard.grad.float(), ref_in.grad[:, rank * Vs:(rank + 1) * Vs].float()
    err["grad_ulp_viol"] = int(((gs - gr).abs() > 2.0 ** -7 * gr.abs() + 1e-6).sum())
    ok = err["logp"] < 2e-5 and err["ent"] < 2e-5 and err["grad"] < 1e-3 and err["grad_ulp_viol"] == 0
    del ref_in
    # timing: sharded fwd+bwd (max over ranks)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    for it in range(13):
        if it == 3:
            dist.barrier(group=group)
            torch.cuda.synchronize()
            ev[0].record()
        shard.grad = None
        lp, ent = mcore_functional.vocab_parallel_logprobs_entropy(shard, labels, 0.9, group=group)
        ((lp * g1).sum() + (ent * g2).sum()).backward()
    ev[1].record()
    torch.cuda.synchronize()
    ms = torch.tensor([ev[0].elapsed_time(ev[1]) / 10, 0.0 if ok else 1.0, err["logp"], err["ent"], err["grad"],
                       float(err["grad_ulp_viol"])], device=dev, dtype=torch.float64)
    if ms.device.type == "cuda" and dist.get_backend(group) == "gloo":
        ms = ms.cpu()
    dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
    return dict(ok=bool(ms[1].item() == 0),
                max_abs_err=dict(logp=float(ms[2]), ent=float(ms[3]), grad=float(ms[4]), grad_ulp_viol=int(ms[5])),
                fwd_bwd_ms=float(ms[0]), tokens_per_s=T / float(ms[0]) * 1e3,
                shard_GBps=6 * Vs * T / float(ms[0]) / 1e6)


def run(dev, rank, world, T=4096, V=151936, group=None):
    """Collective over the (already initialised) default group; returns the result dict on every
    rank ("ok" is the AND over ranks and over both transports). The exchange of the [T, 8] records is
    checked and timed twice: through the group's all-gather (NCCL, or gloo when the ranks share a
    GPU) and through peer stores into IPC-mapped buffers (mcore_functional.PeerExchange).
    bench.py --gpus N calls this for its `parity.tp` entry."""
    group = dist.group.WORLD if group is None else group
    assert V % world == 0
    gen = torch.Generator(device=dev).manual_seed(7)
    full = (torch.randn(T, V, generator=gen, device=dev) * 2).bfloat16()
    labels = torch.randint(0, V, (T,), generator=gen, device=dev)
    g1 = torch.randn(T, generator=gen, device=dev)
    g2 = torch.randn(T, generator=gen, device=dev) * 0.1
    res = dict(tp=world, T=T, V=V)
    res["all_gather"] = _one(dev, rank, world, T, V, group, full, labels, g1, g2)
    # the peer-memory transport is optional (it needs CUDA IPC between the ranks): when it cannot be
    # SET UP the entry says so and parity rests on the all-gather path; once set up, a wrong result
    # or a timed-out wait fails the check like any other
    ex = None
    try:
        ex = mcore_functional.enable_peer_exchange(group, max_tokens=max(T, 256))
    except Exception as e:
        res["peer_stores"] = dict(available=False, error=repr(e)[:300])
    if ex is not None:
        res["peer_stores"] = _one(dev, rank, world, T, V, group, full, labels, g1, g2)
        try:
            ex.check()
        except Exception as e:
            res["peer_stores"].update(ok=False, error=repr(e)[:300])
        mcore_functional.disable_peer_exchange(group)
    res["ok"] = bool(res["all_gather"]["ok"] and res["peer_stores"].get("ok", not res["peer_stores"].get("available", True)))
    # headline numbers of the entry = the faster transport
    best = min((r for r in (res["all_gather"], res["peer_stores"]) if "fwd_bwd_ms" in r), key=lambda r: r["fwd_bwd_ms"])
    res.update(max_abs_err=best["max_abs_err"], fwd_bwd_ms=best["fwd_bwd_ms"], tokens_per_s=best["tokens_per_s"],
               shard_GBps=best["shard_GBps"])
    return res


def main():
    """torchrun entry. With B200RL_ONE_GPU=1 every rank uses cuda:0 and the group is gloo (NCCL
    refuses two ranks on one device): the shard kernels, the exchange and the combine still run
    on the GPU, only the transport of the [T, 8] records differs."""
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    one_gpu = os.environ.get("B200RL_ONE_GPU") == "1"
    dev = torch.device("cuda", 0 if one_gpu else int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    if one_gpu:
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=dev)
    res = run(dev, rank, world, T=int(sys.argv[1]) if len(sys.argv) > 1 else 4096)
    res["transport"] = "gloo (two ranks on one GPU)" if one_gpu else "nccl"
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()
    sys.exit(0 if res["ok"] else 1)


if __name__ == "__main__":
    main()
