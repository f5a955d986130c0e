"""Fused slice -> (NVLink) -> set between two ranks: run under torchrun with >= 2 GPUs.
Rank 0 owns the source flat buffer. Rank 1 (a) does what the reference does - rank 0 slices into a
staging buffer, NCCL broadcast, rank 1 set_intervals - and (b) opens rank 0's buffer as peer
memory and runs move_intervals once. Both results must be bit-identical; both are timed with CUDA
events on rank 1 (the slice on rank 0 is timed there and added to (a)). One JSON line on rank 0."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_vllm_b200 import dist as bdist  # noqa: E402
from areal_vllm_b200 import flatten_param as FP  # noqa: E402


def tp_reshard_lists(scale=1.0):
    """Qwen2.5-7B TP 1->2 style mix: per layer two row-linear weights become one interval per
    matrix row (1,792 / 9,472 elements), everything else long merged intervals."""
    hidden, inter, layers = 3584, 18944, int(28 * scale)
    s_iv, pos = [], 0
    for _ in range(layers):
        for rows, width in ((hidden, hidden), (hidden, inter)):   # attn.c_proj, down_proj
            half = width // 2
            starts = pos + np.arange(rows, dtype=np.int64) * width
            s_iv.append(np.stack([starts, starts + half], 1))
            pos += rows * width
        long_len = (hidden * hidden * 3 + hidden * inter * 2) // 2
        s_iv.append(np.array([[pos, pos + long_len]], dtype=np.int64))
        pos += long_len * 2
    s_iv = np.concatenate(s_iv)
    total = int((s_iv[:, 1] - s_iv[:, 0]).sum())
    # receiver: contiguous per-layer blocks (merged), i.e. few long intervals with small gaps
    d_iv, dpos, left = [], 0, total
    per = total // layers
    for k in range(layers):
        ln = per if k < layers - 1 else left
        d_iv.append((dpos + 64, dpos + 64 + ln))
        dpos += 64 + ln
        left -= ln
    return s_iv, np.array(d_iv, dtype=np.int64), pos, dpos + 64, total


def run(dev, rank, world, scale=0.5, iters=4):
    """Collective over the default group (nccl, or gloo when the ranks share one GPU). Rank 0 owns
    the source; EVERY other rank does (a) the reference's sequence - slice on rank 0, broadcast,
    set - and (b) one move_intervals reading rank 0's buffer through CUDA IPC, and compares the
    two destinations bit for bit. Returns the result dict on every rank."""
    s_iv, d_iv, src_size, dst_size, total = tp_reshard_lists(scale)
    host_bcast = dist.get_backend() == "gloo"
    gen = torch.Generator(device=dev).manual_seed(1)
    src = None
    if rank == 0:
        src = torch.randint(-32768, 32767, (src_size,), dtype=torch.int16, device=dev, generator=gen).view(torch.bfloat16)
    s_c = torch.from_numpy(s_iv).to(dev)
    d_c = torch.from_numpy(d_iv).to(dev)
    peer_src = bdist.open_peer_buffer(src, 0)

    def bcast(buf):
        if host_bcast:  # gloo: stage through the host (two ranks on one GPU; timing is not the point)
            h = buf.view(torch.uint8).cpu()  # gloo moves bytes, not int16/bf16
            dist.broadcast(h, src=0)
            buf.view(torch.uint8).copy_(h)
        else:
            dist.broadcast(buf, src=0)

    ev = lambda: torch.cuda.Event(enable_timing=True)
    # (a) reference sequence
    dst_a = torch.zeros(dst_size, dtype=torch.bfloat16, device=dev) if rank else None
    t_a = []
    for it in range(iters):
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = ev(), ev()
        e0.record()
        if rank == 0:
            buf = FP.slice_intervals(src, intervals_cuda=s_c, output_size=total)
        else:
            buf = torch.empty(total, dtype=torch.bfloat16, device=dev)
        bcast(buf)
        if rank:
            FP.set_intervals(buf, dst_a, intervals_cuda=d_c)
        e1.record(); torch.cuda.synchronize()
        t_a.append(e0.elapsed_time(e1))
    del buf
    # (b) fused peer move on every consumer rank at once (all pull from rank 0's HBM)
    dst_b = torch.zeros(dst_size, dtype=torch.bfloat16, device=dev) if rank else None
    t_b = []
    for it in range(iters):
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = ev(), ev()
        e0.record()
        if rank:
            FP.move_intervals(peer_src, dst_b, s_c, d_c, total)
        e1.record(); torch.cuda.synchronize()
        t_b.append(e0.elapsed_time(e1))
    dist.barrier()
    ok = torch.ones(1)
    if rank:
        same = torch.equal(dst_a.view(torch.int16), dst_b.view(torch.int16))
        moved = int((dst_b.view(torch.int16) != 0).sum()) > 0.9 * total  # random payload: ~all non-zero
        ok[0] = float(same and moved)
    ta = torch.tensor([min(t_a[1:]), min(t_b[1:])])
    if not host_bcast:
        ok, ta = ok.to(dev), ta.to(dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    dist.all_reduce(ta, op=dist.ReduceOp.MAX)
    if rank:
        peer_src.close()
    dist.barrier()
    gb = total * 2 / 1e9
    return dict(ok=bool(ok.item() == 1.0), world=world, consumers=world - 1, copied_GB=gb, n_src=len(s_iv),
                n_dst=len(d_iv), ms_slice_bcast_set=float(ta[0]), ms_fused_peer_move=float(ta[1]),
                fused_GBps_per_consumer=gb / float(ta[1]) * 1e3, staged_GBps=gb / float(ta[0]) * 1e3)


def main():
    """torchrun entry; B200RL_ONE_GPU=1 puts every rank on cuda:0 with a gloo group (the IPC
    mapping and the kernel are the same, the bytes just do not cross NVLink)."""
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    one_gpu = os.environ.get("B200RL_ONE_GPU") == "1"
    dev = torch.device("cuda", 0 if one_gpu else int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    if one_gpu:
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=dev)
    res = run(dev, rank, world, float(sys.argv[1]) if len(sys.argv) > 1 else 0.5)
    res["transport"] = "gloo + CUDA IPC on one GPU" if one_gpu else "nccl + CUDA IPC over NVLink"
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()
    sys.exit(0 if res["ok"] else 1)


if __name__ == "__main__":
    main()
