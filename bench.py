#!/usr/bin/env python
"""Headline benchmark of the b200rl hot path (contract: see the task statement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200rl|reference]

Workload = BASELINE.json configs[1]: GSM8K GRPO step shape for Qwen2-1.5B-Instruct
(V = 151,936): 512 sequences <= 1k tokens in groups of 8, packed into micro-batches of
<= 10,240 tokens (examples/math/gsm8k_grpo.yaml:41). One "step" = one pass of the hot path over
the whole batch: group-level advantage normalisation, then per micro-batch
grpo_loss_fn(logits) [log-prob + entropy + row max/min gather, clipped decoupled-PPO loss] and
loss.backward() down to dlogits. Metric: tokens/s (whole job, all ranks).

  value  inputs resident in HBM when the timed region starts (4 rotating 3.1 GB logits buffers,
         each far larger than the 126 MB L2, so nothing is re-read from cache).
  e2e    the same step through the same public API but every micro-batch's logits/labels/
         per-token inputs start in pinned HOST memory (H2D inside the timed region, overlapped on
         a copy stream) and the loss is read back to the host.
  roofline  the dominant kernel (logprob backward: 4*V bytes/token) timed with CUDA events around
         each of its launches inside the timed region, against MEASURED_PEAKS.json.
  cpu_baseline  the same CPU leg as --impl reference on the same bounded sample (4,096 tokens).

--impl reference times the reference's own CPU implementation of the path with all host threads:
the reference's own functions when a reference tree is on the box (/root/reference, or the
pip-installed copy under baseline/_ref that travels with the snapshot), else the oracle port.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

V = 151936
N_SEQ, GROUP, MAX_MB_TOKENS = 512, 8, 10240
TEMPERATURE, EPS_CLIP, CAP = 1.0, 0.4, 5.0
N_LOGITS_BUFFERS = 4
METRIC = "tokens/s: logprob+PPO loss fwd/bwd @V=152k (GRPO step, Qwen2-1.5B shape)"
WORKLOAD = "cfg2: GSM8K GRPO step, Qwen2-1.5B V=151936, 512 seqs<=1k tok, group 8, mb<=10240 tok"


# ------------------------------------------------------------------------------------------
# synthetic batch (SURVEY §8d cfg2 recipe)
# ------------------------------------------------------------------------------------------
def make_batch(seed: int):
    rng = np.random.default_rng(seed)
    seqlens, resp_starts = [], []
    for _ in range(N_SEQ // GROUP):
        p = int(rng.integers(64, 257))
        for _ in range(GROUP):
            r = int(rng.integers(16, 1024 - p + 1))
            seqlens.append(p + r)
            resp_starts.append(p)
    # greedy packing of whole sequences into micro-batches of <= MAX_MB_TOKENS tokens
    mbs, cur, cur_tok = [], [], 0
    for i, n in enumerate(seqlens):
        if cur_tok + n > MAX_MB_TOKENS and cur:
            mbs.append(cur)
            cur, cur_tok = [], 0
        cur.append(i)
        cur_tok += n
    if cur:
        mbs.append(cur)
    rewards = rng.integers(0, 2, N_SEQ).astype(np.float32)
    return seqlens, resp_starts, mbs, rewards


def mb_host_tensors(seqlens, resp_starts, seqs, adv_seq, seed):
    """Per-token inputs of one packed micro-batch (CPU tensors)."""
    gen = torch.Generator().manual_seed(seed)
    T = sum(seqlens[i] for i in seqs)
    input_ids = torch.randint(0, V, (T,), generator=gen)
    loss_mask = torch.zeros(T)
    adv = torch.zeros(T)
    off = 0
    for i in seqs:
        n, p = seqlens[i], resp_starts[i]
        loss_mask[off + p - 1: off + n - 1] = 1.0  # response positions rolled by -1
        adv[off: off + n] = adv_seq[i]
        off += n
    logp_ref = -torch.rand(T, generator=gen) * 4
    return dict(
        input_ids=input_ids,
        loss_mask=loss_mask,
        advantages=adv,
        logprobs=logp_ref + 0.05 * torch.randn(T, generator=gen),   # behaviour policy
        prox_logp=logp_ref + 0.02 * torch.randn(T, generator=gen),  # proximal policy
    )


def fill_logits(buf: torch.Tensor, labels_even_peak: torch.Tensor, seed: int):
    """(randn*2).bf16 with +6 on the label of even rows, generated on the device in slabs."""
    gen = torch.Generator(device=buf.device).manual_seed(seed)
    T = buf.shape[0]
    for lo in range(0, T, 2048):
        hi = min(lo + 2048, T)
        buf[lo:hi] = (torch.randn(hi - lo, V, generator=gen, device=buf.device) * 2).bfloat16()
    rows = torch.arange(0, T, 2, device=buf.device)
    buf[rows, labels_even_peak[rows]] += 6.0


# ------------------------------------------------------------------------------------------
# host placement: run each rank on the CPUs of its GPU's NUMA node, so that the pinned staging
# buffers of the e2e leg are first-touched there and the H2D copies do not cross the socket link
# ------------------------------------------------------------------------------------------
def bind_to_gpu_numa(local: int):
    """-> {"numa_node": n, "cpus": count, "bdf": "..."}; never raises (a box without the sysfs
    entries, or a single-node box, just reports what it found)."""
    info = {"numa_node": None, "cpus": None, "bdf": None}
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bdf = pynvml.nvmlDeviceGetPciInfo(h).busId
        bdf = bdf.decode() if isinstance(bdf, bytes) else bdf
        bdf = bdf.lower()
        if len(bdf.split(":")[0]) == 8:  # nvml prints an 8-digit domain, sysfs a 4-digit one
            bdf = bdf[4:]
        info["bdf"] = bdf
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        if node < 0:
            return info
        info["numa_node"] = node
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["cpus"] = len(cpus)
    except Exception as e:
        info["error"] = repr(e)[:120]
    return info


# ------------------------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md)
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nme, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------
# the reference arm / CPU baseline: the reference's own functions on the host cores
# ------------------------------------------------------------------------------------------
CPU_SAMPLE_TOKENS = 4096  # BASELINE.md §4 / SURVEY §8d: eager fp32 fwd+bwd keeps ~2.7 MB per token alive


class CpuArm:
    """One pass of the hot path over a bounded sample of cfg2 on the host cores, eager fp32 torch
    with all threads. kind "reference": the reference's OWN functions (areal/utils/functional.py:
    10-188 gather_logprobs_entropy + ppo_actor_loss_fn, AdvNorm of areal/engine/ppo/actor.py:370-604)
    loaded from /root/reference or baseline/_ref by oracle/ref_loader.py (the @torch.compile
    decorators are bypassed via _torchdynamo_orig_callable: Inductor's CPU backend does not build
    here, and eager IS the reference's fp32 torch path). kind "port": oracle/torch_ref.py, the
    line-by-line restatement pinned by tests/golden, when no reference tree is on the box."""

    def __init__(self, sample_tokens: int = CPU_SAMPLE_TOKENS):
        from oracle import ref_loader, torch_ref

        torch.set_num_threads(os.cpu_count() or 1)
        self.T = sample_tokens
        self.kind, self.F, self.adv_norm = "port", None, None
        if ref_loader.available():
            try:
                self.F = ref_loader.areal_functional()

                class Cfg:
                    mean_level, std_level, group_size = "group", "group", GROUP

                self.adv_norm = ref_loader.adv_norm_class()(Cfg)
                self.kind = "reference"
                self.where = ref_loader.REF_ROOT
            except Exception as e:  # a broken tree must not cost the line
                self.F, self.why_port = None, repr(e)[:200]
        self.torch_ref = torch_ref
        gen = torch.Generator().manual_seed(0)
        T = self.T
        self.logits = (torch.randn(T, V, generator=gen) * 2).bfloat16()
        self.ids = torch.randint(0, V, (T,), generator=gen)
        self.mask = torch.rand(T, generator=gen) < 0.7
        lp0 = -torch.rand(T, generator=gen) * 4
        self.old = lp0 + 0.05 * torch.randn(T, generator=gen)
        self.prox = lp0 + 0.02 * torch.randn(T, generator=gen)
        self.adv_rows = torch.randn(((T + 63) // 64 + 7) // 8 * 8, 64, generator=gen)

    def step(self, tokens: int | None = None) -> float:
        T = tokens or self.T
        logits = self.logits[:T].detach().requires_grad_(True)
        ids, mask, old, prox = self.ids[:T], self.mask[:T], self.old[:T], self.prox[:T]
        ones = torch.ones_like(self.adv_rows)
        t0 = time.perf_counter()
        if self.F is not None:
            adv = self.adv_norm(self.adv_rows, ones).reshape(-1)[:T]
            lp, ent = self.F.gather_logprobs_entropy(logits, torch.roll(ids, -1), TEMPERATURE)
            loss, _ = self.F.ppo_actor_loss_fn(logprobs=lp, old_logprobs=old, advantages=adv, eps_clip=EPS_CLIP,
                                               loss_mask=mask, c_clip=None, proximal_logprobs=prox,
                                               behav_imp_weight_cap=CAP)
        else:
            adv = self.torch_ref.adv_norm(self.adv_rows, ones, "group", "group", GROUP).reshape(-1)[:T]
            lp, ent = self.torch_ref.logprobs_entropy(logits, torch.roll(ids, -1), TEMPERATURE)
            loss, _ = self.torch_ref.ppo_actor_loss(lp, prox, old, adv, EPS_CLIP, mask, None, None, CAP)
        _ = logits.detach().min(-1).values.float(), logits.detach().max(-1).values.float()  # actor.py:351-352
        loss.backward()
        return time.perf_counter() - t0

    def describe(self, n_runs: str) -> str:
        what = (f"the reference's own areal/utils/functional.py + AdvNorm from {self.where}" if self.kind == "reference"
                else "oracle/torch_ref.py port (no reference tree on this box)")
        return (f"{self.T}-token sample of cfg2 (V={V}), eager fp32 torch fwd+bwd on {os.cpu_count()} host "
                f"threads, {what}, {n_runs}")


def run_cpu_baseline(budget_s: float = 20.0):
    arm = CpuArm()
    arm.step(256)  # warm-up (thread pool, allocator)
    times = []
    t_start = time.perf_counter()
    while not times or (time.perf_counter() - t_start < budget_s and len(times) < 8):
        times.append(arm.step())
    return {"value": arm.T / min(times), "unit": "tokens/s", "cores": os.cpu_count() or 1, "kind": arm.kind,
            "sample": arm.describe(f"best of {len(times)}")}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm()
    for _ in range(max(args.warmup, 1)):
        arm.step(256)
    dt = 0.0
    for k in range(args.steps):
        dt += arm.step()  # compute time only; the (fixed) sample is generated once, outside
    val = arm.T * args.steps / dt
    out = {
        "metric": METRIC, "value": val, "unit": "tokens/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "reference",
        "config": {"workload": WORKLOAD, "note": "CPU path of the reference; each step = a "
                   f"{arm.T}-token sample of the workload (warm-up steps use 256 tokens)"},
        "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": os.cpu_count() or 1, "kind": arm.kind,
                         "sample": arm.describe(f"{args.steps} timed steps")},
        "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)



# ------------------------------------------------------------------------------------------
# secondary metrics of the same path (BASELINE.json: "GAE tokens/s", interval_op GB/s)
# ------------------------------------------------------------------------------------------
def run_extras(dev, peak, compiled_baseline=False):
    import areal_vllm_b200 as pkg
    from areal_vllm_b200 import _lib
    from oracle import cpu_ref

    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device=dev)

    def timeit(fn, iters=10, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(iters):
            flush.zero_()  # L2 flush between timed iterations (inputs below can be < L2)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts))

    out = {"gae": [], "interval_op": [], "torch_gpu_path": None}
    # The reference's own GPU path for H1+H2 — the same torch ops (areal/utils/functional.py:10-26,
    # 124-188), eager, in 1024-row chunks like its wrappers — timed on this B200: the kernels to beat.
    try:
        import torch.nn.functional as F

        Tt = 4096
        gen = torch.Generator(device=dev).manual_seed(11)
        lg = (torch.randn(Tt, V, generator=gen, device=dev) * 2).bfloat16()
        ids = torch.randint(0, V, (Tt,), generator=gen, device=dev)
        lp0 = -torch.rand(Tt, generator=gen, device=dev) * 4
        old, prox = lp0 + 0.05, lp0 + 0.02
        advt = torch.randn(Tt, generator=gen, device=dev)
        msk = torch.rand(Tt, generator=gen, device=dev) < 0.7

        def torch_step():
            x = lg.detach().requires_grad_(True)
            lps, ents = [], []
            for lo in range(0, Tt, 1024):
                la = F.log_softmax(x[lo:lo + 1024].float() / TEMPERATURE, dim=-1)
                ents.append(-(la.exp() * la).sum(-1))
                lps.append(la.gather(-1, ids[lo:lo + 1024, None]).squeeze(-1))
            lp = torch.cat(lps)
            n = msk.count_nonzero() or 1
            ratio = torch.where(msk, torch.exp(lp - prox), 0)
            pg = torch.max(-advt * ratio, -advt * ratio.clamp(1 - EPS_CLIP, 1 + EPS_CLIP))
            bw = (prox - old).exp()
            bm = (bw <= CAP) & msk
            loss = torch.where(msk, pg * torch.where(bm, bw, 0.0), 0).sum() / n
            _ = x.detach().min(-1).values, x.detach().max(-1).values
            loss.backward()

        ms = timeit(torch_step, iters=5, warm=2)
        out["torch_gpu_path"] = {"what": "eager torch restatement of the reference GPU path (log_softmax/gather/"
                                         "entropy in 1024-row chunks, PPO loss, min/max, backward) on this GPU",
                                 "tokens": Tt, "ms": ms, "tokens_per_s": Tt / (ms * 1e-3)}
        if compiled_baseline:
            # the reference decorates the per-chunk function with @torch.compile (functional.py:18-26)
            try:
                @torch.compile
                def chunk_fn(xc, yc, temperature: float):
                    la = F.log_softmax(xc.float() / temperature, dim=-1)
                    ent = -(la.exp() * la).sum(-1)
                    return la.gather(-1, yc.unsqueeze(-1)).squeeze(-1), ent

                def compiled_step():
                    x = lg.detach().requires_grad_(True)
                    lps = [chunk_fn(x[lo:lo + 1024], ids[lo:lo + 1024], TEMPERATURE)[0] for lo in range(0, Tt, 1024)]
                    lp = torch.cat(lps)
                    n = msk.count_nonzero() or 1
                    ratio = torch.where(msk, torch.exp(lp - prox), 0)
                    pg = torch.max(-advt * ratio, -advt * ratio.clamp(1 - EPS_CLIP, 1 + EPS_CLIP))
                    bw = (prox - old).exp()
                    bm = (bw <= CAP) & msk
                    loss = torch.where(msk, pg * torch.where(bm, bw, 0.0), 0).sum() / n
                    _ = x.detach().min(-1).values, x.detach().max(-1).values
                    loss.backward()

                ms_c = timeit(compiled_step, iters=5, warm=3)
                out["torch_gpu_path"]["torch_compile"] = {
                    "what": "same, per-chunk function under torch.compile (Inductor/Triton) as in the reference",
                    "ms": ms_c, "tokens_per_s": Tt / (ms_c * 1e-3)}
            except Exception as e:
                out["torch_gpu_path"]["torch_compile"] = {"error": repr(e)[:300]}
        del lg
        torch.cuda.empty_cache()
    except Exception as e:  # an OOM or similar here must not cost the headline line
        out["torch_gpu_path"] = {"error": repr(e)[:200]}
    # The two other ways the reference enters H1: forward only, once per batch, for
    # the proximal log-probs (`calc_logprobs`, areal/engine/ppo/actor.py:55-58: 2*V B/token), and
    # fp32 logits as the Megatron engine hands them over (megatron_engine.py:510-521: 4*V forward,
    # 8*V backward). Same T as a cfg2 micro-batch; the inputs are far larger than L2.
    try:
        Th = 10240
        gen = torch.Generator(device=dev).manual_seed(12)
        variants = []
        for dt, esz in ((torch.bfloat16, 2), (torch.float32, 4)):
            xh = torch.empty(Th, V, dtype=dt, device=dev)
            for lo_ in range(0, Th, 2048):
                xh[lo_:lo_ + 2048] = (torch.randn(2048, V, generator=gen, device=dev) * 2).to(dt)
            yh = torch.randint(0, V, (Th,), generator=gen, device=dev)
            gl = torch.randn(Th, generator=gen, device=dev)
            ms_f = timeit(lambda: pkg.functional.gather_logprobs(xh, yh, TEMPERATURE), iters=10)
            ms_fe = timeit(lambda: pkg.functional.gather_logprobs_entropy(xh, yh, TEMPERATURE), iters=10)
            _, _, lse_h, _, _ = pkg.ops.logprob_entropy_fwd(xh, yh, TEMPERATURE, False, False)
            ms_b = timeit(lambda: pkg.ops.logprob_entropy_bwd(xh, yh, lse_h, lse_h, gl, None, TEMPERATURE), iters=10)
            for what, ms_, nbytes in (("forward only: gather_logprobs (compute_logp)", ms_f, esz * V),
                                      ("forward: gather_logprobs_entropy", ms_fe, esz * V),
                                      ("backward: dlogits", ms_b, 2 * esz * V)):
                gbps = nbytes * Th / (ms_ * 1e-3) / 1e9
                variants.append({"logits": str(dt).replace("torch.", ""), "call": what, "tokens": Th, "us": ms_ * 1e3,
                                 "tokens_per_s": Th / (ms_ * 1e-3), "algorithmic_bytes_per_token": nbytes,
                                 "GBps": gbps, "frac_of_measured_peak": gbps / peak})
            del xh
            torch.cuda.empty_cache()
        out["h1_call_site_variants"] = variants
    except Exception as e:
        out["h1_call_site_variants"] = {"error": repr(e)[:200]}
    for name, B, lo, hi, seed in (("cfg1: 256 seqs, len 64-2048", 256, 64, 2048, 1),
                                  ("cfg1 length law x256: 65536 seqs", 65536, 64, 2048, 2)):
        rng = np.random.default_rng(seed)
        lens = rng.integers(lo, hi + 1, B)
        total = int(lens.sum())
        r_np = rng.standard_normal(total).astype(np.float32)
        v_np = rng.standard_normal(total + B).astype(np.float32)
        cu_np = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        r, v, cu = (torch.from_numpy(a).to(dev) for a in (r_np, v_np, cu_np))
        b = torch.ones(B, dtype=torch.bool, device=dev)
        for gamma, lam in ((0.99, 0.95), (1.0, 1.0)):
            ms = timeit(lambda: pkg.ppo_functional.cugae1d_nolp_misalign_func(r, v, cu, b, gamma, lam))
            rec = {"workload": name, "gamma": gamma, "lam": lam, "tokens": total,
                   "kernel": "gae_pieces_kernel (tiled scan)" if gamma * lam <= 0.99 else
                   ("gae_serial_relay_kernel" if B >= 1024 else
                    "gae_serial_team_kernel" if B <= 4 * 148 and total >= 512 * B else "gae_serial_kernel"),
                   "us": ms * 1e3, "tokens_per_s": total / (ms * 1e-3),
                   "hbm_frac_of_measured_peak": 16.0 * total / (ms * 1e-3) / 1e9 / peak}
            if B == 256:  # CPU baseline: C port of pygae1d_nolp_misalign, one thread (the python
                t0 = time.perf_counter()  # reference is a scalar loop and cannot use more)
                cpu_ref.gae_py(r_np, v_np, cu_np, np.ones(B, np.float32), gamma, lam)
                rec["cpu_port_tokens_per_s"] = total / (time.perf_counter() - t0)
            out["gae"].append(rec)
    # interval_op, BASELINE config 5: 7B bf16 flat buffer, 1e5 intervals (reference test's generator)
    size, n = 7_615_616_512, 100000
    free, _ = torch.cuda.mem_get_info()
    if free > 26 * 2**30:
        s_ = size // n
        iv = np.array([(i * s_, i * s_ + s_ // 2) for i in range(n)], dtype=np.int64)
        np.random.default_rng(5).shuffle(iv)
        src = torch.empty(size, dtype=torch.bfloat16, device=dev)
        src.view(torch.int16).random_()
        ivc = torch.from_numpy(iv).to(dev)
        total = int((iv[:, 1] - iv[:, 0]).sum())
        packed = pkg.flatten_param.slice_intervals(src, intervals_cuda=ivc, output_size=total)
        ms_s = timeit(lambda: pkg.flatten_param.slice_intervals(src, intervals_cuda=ivc, output_size=total), iters=5)
        ms_t = timeit(lambda: pkg.flatten_param.set_intervals(packed, src, intervals_cuda=ivc), iters=5)
        gb = 4.0 * total / 1e9
        # fused slice->set into a second flat buffer (different cut: 1/3 as many, longer intervals)
        rec_move = {}
        if free > 44 * 2**30:
            dst = torch.empty(size, dtype=torch.bfloat16, device=dev)
            n_d = n // 3
            per = total // n_d
            d_iv = np.array([(i * (size // n_d), i * (size // n_d) + (per if i < n_d - 1 else total - per * (n_d - 1)))
                             for i in range(n_d)], dtype=np.int64)
            dvc = torch.from_numpy(d_iv).to(dev)
            ms_m = timeit(lambda: pkg.flatten_param.move_intervals(src, dst, ivc, dvc, total), iters=5)
            rec_move = {"move_us": ms_m * 1e3, "move_GBps": gb / (ms_m * 1e-3),
                        "move_frac_of_measured_peak": gb / (ms_m * 1e-3) / peak,
                        "move_note": "slice+set fused (no staging buffer): same bytes as ONE of the two passes"}
            del dst
        out["interval_op"].append({
            "workload": "cfg5: 7,615,616,512-element bf16 buffer, 1e5 shuffled intervals (half the buffer)",
            "copied_GB": 2.0 * total / 1e9, "slice_us": ms_s * 1e3, "slice_GBps": gb / (ms_s * 1e-3),
            "slice_frac_of_measured_peak": gb / (ms_s * 1e-3) / peak, "set_us": ms_t * 1e3,
            "set_GBps": gb / (ms_t * 1e-3), "set_frac_of_measured_peak": gb / (ms_t * 1e-3) / peak, **rec_move})
        del src, packed
    return out

def roofline_gae(extras, peak, peak_src):
    """BASELINE.json's metric is "...; GAE tokens/s": the GAE kernels' own roofline entries (16
    algorithmic bytes per token: r, v in; adv, ret out), one per (gamma, lam) of configs[0], at the
    size where a roofline means something (65,536 sequences of the cfg1 length law, 1.1 GB) and
    at cfg1 itself (256 sequences = 4.4 MB: launch-latency bound), with the CPU baseline beside it."""
    if not extras or not extras.get("gae"):
        return None
    out = []
    for g in extras["gae"]:
        gbps = 16.0 * g["tokens"] / (g["us"] * 1e-6) / 1e9
        out.append({"kernel": g["kernel"], "workload": g["workload"], "gamma": g["gamma"], "lam": g["lam"],
                    "bound": "hbm", "achieved": gbps, "peak": peak, "unit": "GB/s", "frac": gbps / peak,
                    "peak_source": peak_src, "algorithmic_bytes_per_token": 16, "tokens_per_s": g["tokens_per_s"],
                    "cpu_baseline_tokens_per_s": g.get("cpu_port_tokens_per_s"),
                    "cpu_baseline_kind": "port (C restatement of pygae1d_nolp_misalign, 1 thread)"
                    if "cpu_port_tokens_per_s" in g else None})
    return out


# ------------------------------------------------------------------------------------------
# BASELINE config 4 (long-CoT, strong scaling): 1,048,576 tokens of 1k-16k-token sequences, whole
# GRPO groups sharded over the ranks by tokens, batch-level advantage normalisation (the one
# collective of the path), 32,768-token micro-batches over a reused V=152,064 logits buffer.
# ------------------------------------------------------------------------------------------
def _gather_var(x: torch.Tensor, world: int):
    """all-gather of per-rank 1-D tensors of different lengths -> list of tensors (every rank)."""
    import torch.distributed as dist

    n = torch.tensor([x.numel()], device=x.device, dtype=torch.int64)
    ns = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(ns, n)
    ns = [int(v) for v in ns]
    pad = torch.zeros(max(ns), dtype=x.dtype, device=x.device)
    pad[: x.numel()] = x
    out = torch.empty(world * max(ns), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, pad)
    return [out[r * max(ns): r * max(ns) + ns[r]] for r in range(world)]


def run_cfg4(dev, rank, world, steps=2, check=True):
    """-> (timing record, parity record or None). The parity record is what a data-parallel run
    has to get right (ref fsdp_engine.py:486-491, 543-547; functional.py:108-115): the all-reduced
    batch moments, the normalised advantages, the summed loss and d loss/d logp of the SHARDED
    step against rank 0 recomputing all of it from the gathered per-token inputs in one piece."""
    import torch.distributed as dist

    import areal_vllm_b200 as pkg
    from areal_vllm_b200 import functional as Fn
    from areal_vllm_b200.actor import grpo_loss_fn

    V4, MB, TOTAL = 152064, 32768, 1_048_576
    CLIP = (0.2, 0.28, 3.0, 5.0)
    rng = np.random.default_rng(4)
    lens = []
    while sum(lens) < TOTAL:
        lens.append(int(rng.integers(1024, 16385)))
    lens[-1] -= sum(lens) - TOTAL
    # token-balanced split at SEQUENCE granularity (areal/utils/data.py:219-243); nothing in the
    # step needs a GRPO group on one rank (batch-level statistics here; group-level ones are
    # additive moments, see dist.sharded_group_advantage_norm)
    mine = pkg.dist.shard_sequences_by_tokens(lens, world)[rank]
    n_tok = sum(lens[i] for i in mine)
    gen = torch.Generator(device=dev).manual_seed(40 + rank)
    adv_raw = torch.randn(n_tok, generator=gen, device=dev)
    mask = torch.rand(n_tok, generator=gen, device=dev) < 0.8
    ids = torch.randint(0, V4, (n_tok,), generator=gen, device=dev)
    lp0 = -torch.rand(n_tok, generator=gen, device=dev) * 4

    def make_buf(r):
        g = torch.Generator(device=dev).manual_seed(400 + r)
        b = torch.empty(MB, V4, dtype=torch.bfloat16, device=dev)
        for lo in range(0, MB, 2048):
            b[lo:lo + 2048] = (torch.randn(2048, V4, generator=g, device=dev) * 2).bfloat16()
        return b

    buf = make_buf(rank)
    bounds = list(range(0, n_tok, MB)) + [n_tok]
    old_lp, prox_lp = lp0 + 0.05, lp0 + 0.02

    def stats_step():
        # batch-level mean/std over ALL ranks + the global loss normaliser: one all-reduce
        if world > 1:
            return pkg.dist.global_advantage_stats(adv_raw, mask)
        mom = pkg.ops.masked_moments(adv_raw, mask.float(), n_tok, None, 1)
        mean, denom = pkg.ops.moments_finalize(mom, False, False, 1e-5)
        return mean, denom, mom[0, 0:1]

    def step():
        mean, denom, total_w = stats_step()
        adv = ((adv_raw * mask - mean.float()) / denom.float()).float()
        tot = torch.zeros((), device=dev)
        for a, b in zip(bounds[:-1], bounds[1:]):
            x = buf[: b - a].detach().requires_grad_(True)
            data = dict(input_ids=ids[a:b], logprobs=old_lp[a:b], advantages=adv[a:b],
                        loss_mask=mask[a:b], prox_logp=prox_lp[a:b])
            loss = grpo_loss_fn(x, data, 1.0, *CLIP, stats={})
            # the engine's scaling: local mask count / global count (fsdp_engine.py:543-547)
            (g,) = torch.autograd.grad(loss * (mask[a:b].sum() / total_w.float()), x)
            tot = tot + loss.detach()
            del g
        return tot

    step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    toks = torch.tensor([float(n_tok), ms], device=dev, dtype=torch.float64)
    if world > 1:
        mx = toks.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(toks)
        max_tok, ms, sum_tok = float(mx[0]), float(mx[1]), float(toks[0])
    else:
        max_tok, sum_tok = float(n_tok), float(n_tok)
    rec = {"workload": "cfg4: 1,048,576 tokens, seqs 1k-16k, sharded by tokens at sequence granularity, "
                       "V=152064, mb 32768 tok on a reused logits buffer, batch-level AdvNorm (1 all-reduce/"
                       "step), eps_clip 0.2/0.28, c_clip 3, cap 5",
           "scaling": "strong", "n_gpus": world, "tokens": sum_tok, "ms_per_step_max_over_ranks": ms,
           "tokens_per_s": sum_tok / (ms * 1e-3), "imbalance_max_over_mean_tokens": max_tok / (sum_tok / world),
           "per_gpu_frac_of_measured_peak_at_6V_on_busiest_rank": None}

    parity = None
    if check and world > 1:
        # ---- untimed: one more sharded step, by pieces, keeping per-token results ----
        mean, denom, total_w = stats_step()
        adv = ((adv_raw * mask - mean.float()) / denom.float()).float()
        logp_l, g_l = [], []
        scaled_loss = torch.zeros((), device=dev, dtype=torch.float64)
        for a, b in zip(bounds[:-1], bounds[1:]):
            x = buf[: b - a].detach().requires_grad_(True)
            lp, _, _, _ = Fn.gather_logprobs_entropy_minmax(x, torch.roll(ids[a:b], -1), 1.0)
            lp.retain_grad()
            loss, _ = Fn.ppo_actor_loss_fn(lp, prox_lp[a:b], old_lp[a:b], adv[a:b], CLIP[0], mask[a:b],
                                           CLIP[1], CLIP[2], CLIP[3])
            w = (mask[a:b].sum() / total_w.float()).reshape(())
            (loss * w).backward()
            scaled_loss += (loss.detach() * w).double()
            logp_l.append(lp.detach())
            g_l.append(lp.grad.detach())
            del x
        logp_mine, g_mine = torch.cat(logp_l), torch.cat(g_l)
        dist.all_reduce(scaled_loss)
        parts = {k: _gather_var(v, world) for k, v in dict(
            adv_raw=adv_raw, mask=mask.float(), ids=ids.double(), lp0=lp0, logp=logp_mine, g=g_mine, adv=adv).items()}
        ok = torch.ones(1, device=dev)
        if rank == 0:
            cat = {k: torch.cat(v) for k, v in parts.items()}
            m_all = cat["mask"] > 0
            mom = pkg.ops.masked_moments(cat["adv_raw"].contiguous(), cat["mask"].contiguous(), cat["mask"].numel(), None, 1)
            mean1, denom1 = pkg.ops.moments_finalize(mom, False, False, 1e-5)
            adv1 = ((cat["adv_raw"] * cat["mask"] - mean1.float()) / denom1.float()).float()
            lp_all = cat["logp"].clone().requires_grad_(True)
            loss1, _ = Fn.ppo_actor_loss_fn(lp_all, cat["lp0"] + 0.02, cat["lp0"] + 0.05, adv1, CLIP[0], m_all,
                                            CLIP[1], CLIP[2], CLIP[3])
            loss1.backward()
            rel = lambda x_, y_: float((x_ - y_).abs().max() / y_.abs().max().clamp_min(1e-30))
            err = {"mean": float((mean - mean1).abs().max()), "denom_rel": rel(denom, denom1),
                   "total_loss_weight": float((total_w.double() - mom[0, 0]).abs().max()),
                   "adv_max_abs": float((cat["adv"] - adv1).abs().max()),
                   "loss_rel": float(abs(scaled_loss.item() - loss1.item()) / max(abs(loss1.item()), 1e-30)),
                   "g_logp_rel": rel(cat["g"], lp_all.grad)}
            # another GPU's log-probs, recomputed here from its (regenerated) logits and gathered ids
            worst = 0.0
            for r in range(1, world):
                rb = make_buf(r)
                ids_r, lp_r = parts["ids"][r].long(), parts["logp"][r]
                for a in range(0, ids_r.numel(), MB):
                    b = min(a + MB, ids_r.numel())
                    lp_here, _, _, _ = Fn.gather_logprobs_entropy_minmax(rb[: b - a], torch.roll(ids_r[a:b], -1), 1.0)
                    worst = max(worst, float((lp_here - lp_r[a:b]).abs().max()))
                del rb
            err["logp_other_gpus_max_abs"] = worst
            good = (err["mean"] <= 1e-5 and err["denom_rel"] <= 1e-5 and err["total_loss_weight"] == 0.0
                    and err["adv_max_abs"] <= 1e-5 and err["loss_rel"] <= 1e-3 and err["g_logp_rel"] <= 1e-3
                    and worst <= 1e-5)
            ok[0] = float(good)
            parity = {"ok": bool(good), "tokens": int(cat["mask"].numel()), "err": err,
                      "bars": "mean/denom/adv 1e-5, loss and d loss/d logp 1e-3 rel, logp 1e-5"}
        dist.broadcast(ok, src=0)
        if rank != 0:
            parity = {"ok": bool(ok.item() == 1.0)}
    del buf
    torch.cuda.empty_cache()
    return rec, parity


def run_group_norm_parity(dev, rank, world):
    """Group-level AdvNorm with the rows of a group on different ranks (sequence-granular
    sharding) vs AdvNorm(group, group) on the whole batch (ref actor.py:412-470), cfg2 shape."""
    import torch.distributed as dist

    import areal_vllm_b200 as pkg
    from areal_vllm_b200.actor import AdvNorm

    seqlens, resp_starts, _, rewards = make_batch(seed=2)
    L = max(seqlens)
    pad_mask = torch.zeros(N_SEQ, L)
    for i, (n, p) in enumerate(zip(seqlens, resp_starts)):
        pad_mask[i, p - 1: n - 1] = 1.0
    gen = torch.Generator().manual_seed(22)
    adv_pad = ((torch.from_numpy((rewards - 0.5) * 10.0)[:, None] + 0.1 * torch.randn(N_SEQ, L, generator=gen))
               * pad_mask).to(dev)
    pad_mask = pad_mask.to(dev)
    mine = torch.tensor(pkg.dist.shard_sequences_by_tokens(seqlens, world)[rank], device=dev, dtype=torch.int64)
    out, = pkg.dist.sharded_group_advantage_norm(adv_pad[mine], pad_mask[mine], mine // GROUP, N_SEQ // GROUP)

    class Cfg:
        mean_level, std_level, group_size = "group", "group", GROUP

    full = AdvNorm(Cfg)(adv_pad, pad_mask)
    err = torch.tensor([float((out - full[mine]).abs().max())], device=dev)
    dist.all_reduce(err, op=dist.ReduceOp.MAX)
    return {"ok": bool(err.item() <= 1e-5), "max_abs_err": float(err.item()), "rows": N_SEQ, "groups": N_SEQ // GROUP}


# ------------------------------------------------------------------------------------------
# the b200rl arm
# ------------------------------------------------------------------------------------------
def main_b200rl(args):
    import torch.distributed as dist

    import areal_vllm_b200 as pkg
    from areal_vllm_b200 import ops
    from areal_vllm_b200.actor import AdvNorm, grpo_loss_fn, grpo_loss_fn_fused

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    numa = bind_to_gpu_numa(local) if world > 1 else {"numa_node": None, "note": "single rank: not bound"}
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    # weak scaling: every rank owns one full GRPO batch (its own seed); no data-path collective,
    # one tiny all-reduce per step for the global loss normaliser (fsdp_engine.py:491)
    seqlens, resp_starts, mbs, rewards = make_batch(seed=2 + rank)
    n_tok = sum(seqlens)
    L = max(seqlens)

    class Cfg:
        mean_level, std_level, group_size = "group", "group", GROUP

    adv_norm = AdvNorm(Cfg)
    # padded [B, L] advantages as compute_advantages leaves them for gamma=lam=1, values=0:
    # the scaled sequence reward on every response position (actor.py:85-147)
    reward_score = torch.from_numpy((rewards - 0.5) * 10.0)
    pad_mask = torch.zeros(N_SEQ, L)
    for i, (n, p) in enumerate(zip(seqlens, resp_starts)):
        pad_mask[i, p - 1: n - 1] = 1.0
    adv_pad = (reward_score[:, None] * pad_mask).to(dev)
    pad_mask = pad_mask.to(dev)

    # normalised advantages decide the per-token inputs; compute once for data generation
    adv_n = adv_norm(adv_pad, pad_mask)
    adv_seq = [float(adv_n[i, resp_starts[i] - 1]) for i in range(N_SEQ)]
    host_mbs = [mb_host_tensors(seqlens, resp_starts, seqs, adv_seq, seed=1000 * rank + k)
                for k, seqs in enumerate(mbs)]
    dev_mbs = [{k: v.to(dev) for k, v in mb.items()} for mb in host_mbs]
    max_T = max(mb["input_ids"].numel() for mb in host_mbs)

    logits_bufs = []
    for b in range(N_LOGITS_BUFFERS):
        buf = torch.empty(max_T, V, dtype=torch.bfloat16, device=dev)
        lab = torch.roll(dev_mbs[b % len(dev_mbs)]["input_ids"], -1)
        lab = torch.cat([lab, lab.new_zeros(max_T - lab.numel())])
        fill_logits(buf, lab, seed=77 + b + 10 * rank)
        logits_bufs.append(buf)

    bwd_events = []
    fwd_events = []

    unit_scale = torch.ones((), device=dev)

    def hot_step_fused(step_idx: int):
        """The same step through the fused extension API (grpo_loss_fn_fused: loss_scale handed
        in up front, forward + loss + dlogits in one kernel, 4*V bytes of HBM traffic per token)."""
        adv_norm(adv_pad, pad_mask)
        total = torch.zeros((), device=dev)
        for k, mb in enumerate(dev_mbs):
            T = mb["input_ids"].numel()
            x = logits_bufs[(k + step_idx) % N_LOGITS_BUFFERS][:T].detach().requires_grad_(True)
            loss = grpo_loss_fn_fused(x, mb, TEMPERATURE, EPS_CLIP, None, None, CAP, loss_scale=unit_scale,
                                      stats={})  # the logged tensors are produced, as in the engine
            (dlogits,) = torch.autograd.grad(loss, x)
            total = total + loss.detach()
            del dlogits
        return total

    def hot_step(step_idx: int, record: bool, src_logits=None):
        """One pass of the hot path over the whole batch; returns the summed loss (device)."""
        adv_norm(adv_pad, pad_mask)  # group mean/std advantage normalisation of the step
        if world > 1:
            w = torch.tensor([float(n_tok)], device=dev, dtype=torch.float64)
            pkg.dist.allreduce_stats(w)
        total = torch.zeros((), device=dev)
        for k, mb in enumerate(dev_mbs):
            T = mb["input_ids"].numel()
            logits = (src_logits(k) if src_logits else logits_bufs[(k + step_idx) % N_LOGITS_BUFFERS])[:T]
            x = logits.detach().requires_grad_(True)
            if record:
                e0, e1, e3 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
                e0.record()
            # stats={}: every tensor the reference registers with stats_tracker (actor.py:325-365) is
            # produced inside the timed region, as it is in the engine
            loss = grpo_loss_fn(x, mb, TEMPERATURE, EPS_CLIP, None, None, CAP, stats={})
            if record:
                e1.record()
            (dlogits,) = torch.autograd.grad(loss, x)
            if record:
                e3.record()
                fwd_events.append((e0, e1, T))
                bwd_events.append((e1, e3, T))
            total = total + loss.detach()
            del dlogits
        return total

    # ---------------- timed region: inputs resident in HBM ----------------
    for w in range(args.warmup):
        hot_step(w, False)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    torch.cuda.synchronize()
    t_ev0, t_ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_ev0.record()
    for s in range(args.steps):
        hot_step(args.warmup + s, True)
    t_ev1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    elapsed_ms = t_ev0.elapsed_time(t_ev1)
    if world > 1:
        dist.barrier()
        t = torch.tensor([elapsed_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
        tok = torch.tensor([float(n_tok)], device=dev, dtype=torch.float64)
        dist.all_reduce(tok)
        total_tok = float(tok.item())
        toks = [torch.zeros_like(tok) for _ in range(world)]
        dist.all_gather(toks, torch.tensor([float(n_tok)], device=dev, dtype=torch.float64))
        tokens_per_rank = [int(t.item()) for t in toks]
    else:
        total_tok = float(n_tok)
        tokens_per_rank = [n_tok]
    value = total_tok * args.steps / (elapsed_ms * 1e-3)

    # per-launch device times of the composite forward and of the backward chain.
    # "bwd" = ppo_loss_bwd (tiny) + logprob_entropy_bwd; "fwd" = logprob fwd + loss fwd + stats.
    bwd_ms = sum(a.elapsed_time(b) for a, b, _ in bwd_events)
    fwd_ms = sum(a.elapsed_time(b) for a, b, _ in fwd_events)
    bwd_bytes = sum(4.0 * V * T for _, _, T in bwd_events)
    fwd_bytes = sum(2.0 * V * T for _, _, T in fwd_events)

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):  # dram__bytes_read+write per [T,V] row from the last ncu --set full capture
        per_row = json.load(open(tpath)).get("logprob_bwd_bf16_kernel_bytes_per_row")
        if per_row:
            traffic = per_row * (sum(T for _, _, T in bwd_events) / max(len(bwd_events), 1))
    achieved = bwd_bytes / (bwd_ms * 1e-3) / 1e9
    roofline = {
        "kernel": "logprob_bwd_kernel<bf16> (tile-per-CTA, loss backward folded in)", "bound": "hbm",
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": traffic, "peak_source": peak_src,
        "algorithmic_bytes_per_launch": bwd_bytes / max(len(bwd_events), 1),
        "algorithmic_bytes_per_token": 4 * V, "launches_timed": len(bwd_events),
        "fwd": {"kernel": "logprob_fwd_kernel<bf16,ENT,MINMAX> + ppo_loss_fwd_kernel",
                "achieved": fwd_bytes / (fwd_ms * 1e-3) / 1e9,
                "frac": fwd_bytes / (fwd_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_token": 2 * V},
        "fwd_bwd_frac_of_peak": (fwd_bytes + bwd_bytes) / ((fwd_ms + bwd_ms) * 1e-3) / 1e9 / peak,
    }

    # ---------------- the fused extension path, same step, same buffers ----------------
    fused = None
    if not args.no_fused:
        for w in range(max(args.warmup, 1)):
            hot_step_fused(w)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for s_ in range(args.steps):
            hot_step_fused(args.warmup + s_)
        f1.record()
        torch.cuda.synchronize()
        fms = f0.elapsed_time(f1)
        if world > 1:
            t = torch.tensor([fms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            fms = float(t.item())
        fval = total_tok * args.steps / (fms * 1e-3)
        fused = {"api": "areal_vllm_b200.actor.grpo_loss_fn_fused(..., loss_scale=...) + backward()",
                 "value": fval, "unit": "tokens/s", "ms_per_step": fms / args.steps,
                 "speedup_vs_separate_path": fval / value,
                 "algorithmic_bytes_per_token": 4 * V,
                 "hbm_GBps_per_gpu_at_4V": fval / world * 4 * V / 1e9,
                 "frac_of_measured_peak_at_4V": fval / world * 4 * V / 1e9 / peak,
                 "launch": ("one row per CTA" if os.environ.get("B200RL_FUSED_PAIR", "1") in ("0",)
                            else "pair mode: clusters of two CTAs sharing two row streams (the library's default; "
                                 "it takes the one-row-per-CTA form itself if no cluster fits on the device)"),
                 "note": "each row is read from HBM once (forward statistics) and again from L2 for the "
                         "backward; needs the loss scale before backward(), see INTEGRATION.md"}

    # ---------------- e2e: host buffers, H2D inside the timed region ----------------
    e2e = None
    if not args.no_e2e:
        n_pin = 2
        pin_logits = [torch.empty(max_T, V, dtype=torch.bfloat16).pin_memory() for _ in range(n_pin)]
        for b in range(n_pin):
            pin_logits[b].copy_(logits_bufs[b])
        pin_mbs = [{k: v.pin_memory() for k, v in mb.items()} for mb in host_mbs]
        stage = [torch.empty(max_T, V, dtype=torch.bfloat16, device=dev) for _ in range(2)]
        copy_stream = torch.cuda.Stream(device=dev)
        loss_host = torch.empty((), dtype=torch.float32).pin_memory()
        h2d_bytes = sum(mb["input_ids"].numel() * (V * 2) + sum(v.numel() * v.element_size() for v in mb.values())
                        for mb in host_mbs)

        def e2e_step():
            main = torch.cuda.current_stream(dev)
            ready = [None, None]
            free = [None, None]

            def prefetch(k):
                T = host_mbs[k]["input_ids"].numel()
                s = k % 2
                with torch.cuda.stream(copy_stream):
                    if free[s] is not None:
                        copy_stream.wait_event(free[s])
                    stage[s][:T].copy_(pin_logits[k % n_pin][:T], non_blocking=True)
                    mbd = {kk: vv.to(dev, non_blocking=True) for kk, vv in pin_mbs[k].items()}
                    ev = torch.cuda.Event()
                    ev.record(copy_stream)
                ready[s] = (ev, mbd)

            adv_norm(adv_pad, pad_mask)
            total = torch.zeros((), device=dev)
            prefetch(0)
            for k in range(len(host_mbs)):
                if k + 1 < len(host_mbs):
                    prefetch(k + 1)
                ev, mbd = ready[k % 2]
                main.wait_event(ev)
                T = mbd["input_ids"].numel()
                x = stage[k % 2][:T].detach().requires_grad_(True)
                loss = grpo_loss_fn(x, mbd, TEMPERATURE, EPS_CLIP, None, None, CAP, stats={})
                (dlogits,) = torch.autograd.grad(loss, x)
                total = total + loss.detach()
                fe = torch.cuda.Event()
                fe.record(main)
                free[k % 2] = fe
                del dlogits
            loss_host.copy_(total, non_blocking=True)
            main.synchronize()
            return float(loss_host)

        e2e_step()  # warm-up
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        n_e2e = max(1, min(args.steps, 2))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_e2e):
            e2e_step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        my_gbps = h2d_bytes * n_e2e / (ms * 1e-3) / 1e9
        per_rank = [{"rank": rank, "h2d_GBps": my_gbps, **numa}]
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            box = [None] * world
            dist.all_gather_object(box, per_rank[0])
            per_rank = box
        e2e = {"value": total_tok * n_e2e / (ms * 1e-3), "unit": "tokens/s",
               "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": 4, "steps": n_e2e,
               "per_rank": per_rank,
               "note": "logits+labels+per-token inputs start in pinned host memory (each rank bound to the "
                       "CPUs of its GPU's NUMA node before pinning); H2D double-buffered on a copy stream; "
                       "loss read back each step; h2d_bytes_per_step is per rank"}
        del pin_logits, stage

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_baseline = run_cpu_baseline()
    extras = None
    if world > 1 and not args.no_extras:
        # GAE shards by sequence too: every rank scans its own 65,536 sequences (weak scaling)
        import areal_vllm_b200 as pkg
        rng = np.random.default_rng(2 + rank)
        lens = rng.integers(64, 2049, 65536)
        tot = int(lens.sum())
        r_ = torch.randn(tot, device=dev)
        v_ = torch.randn(tot + 65536, device=dev)
        cu_ = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
        b_ = torch.ones(65536, dtype=torch.bool, device=dev)
        res = {}
        for gamma, lam in ((0.99, 0.95), (1.0, 1.0)):
            for _ in range(3):
                pkg.ppo_functional.cugae1d_nolp_misalign_func(r_, v_, cu_, b_, gamma, lam)
            torch.cuda.synchronize()
            dist.barrier()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record()
            for _ in range(10):
                pkg.ppo_functional.cugae1d_nolp_misalign_func(r_, v_, cu_, b_, gamma, lam)
            g1.record()
            torch.cuda.synchronize()
            tm = torch.tensor([g0.elapsed_time(g1) / 10, float(tot)], device=dev, dtype=torch.float64)
            tmax = tm.clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(tm, op=dist.ReduceOp.SUM)
            res[f"gamma={gamma},lam={lam}"] = {"tokens_all_ranks": float(tm[1]), "ms_max_over_ranks": float(tmax[0]),
                                               "tokens_per_s": float(tm[1]) / (float(tmax[0]) * 1e-3)}
        extras = {"gae_weak_scaling_65536_seqs_per_rank": res}
        del r_, v_, cu_, b_
    cfg4 = None
    parity = None
    if not args.no_extras:  # every rank takes part (one all-reduce per step)
        del logits_bufs[:]
        torch.cuda.empty_cache()
        try:
            cfg4, sharded = run_cfg4(dev, rank, world, check=not args.no_parity)
            cfg4["per_gpu_frac_of_measured_peak_at_6V_on_busiest_rank"] = (
                cfg4["imbalance_max_over_mean_tokens"] * cfg4["tokens"] / world * 6 * 152064
                / (cfg4["ms_per_step_max_over_ranks"] * 1e-3) / 1e9 / peak)
        except Exception as e:  # must not cost the headline line
            cfg4, sharded = {"error": repr(e)[:300]}, {"ok": False, "error": repr(e)[:300]}
        if world > 1 and not args.no_parity:
            # ---- N >= 2: correctness of every cross-GPU path, after the timed regions ----
            # sharded_step: the data-parallel arithmetic of the cfg4 step (see run_cfg4);
            # group_norm: group-level AdvNorm with a group's rows on different ranks;
            # tp: vocabulary-parallel log-prob/entropy over NCCL with tp = world vs the whole-
            #     vocabulary kernel (ref mcore/functional.py:8-52);
            # peer_move: move_intervals reading rank 0's flat buffer over NVLink vs slice ->
            #     NCCL broadcast -> set (ref real_llm_api.py:721-805), bit-exact.
            sys.path.insert(0, os.path.join(ROOT, "scripts"))
            import p2p_move_check
            import tp_check

            details = {"sharded_step": sharded}
            for name, fn in (("group_norm", lambda: run_group_norm_parity(dev, rank, world)),
                             ("tp", lambda: tp_check.run(dev, rank, world, T=2048)),
                             ("peer_move", lambda: p2p_move_check.run(dev, rank, world, scale=0.25))):
                try:
                    details[name] = fn()
                except Exception as e:
                    details[name] = {"ok": False, "error": repr(e)[:300]}
                torch.cuda.empty_cache()
            parity = {k: bool(v and v.get("ok")) for k, v in details.items()}
            parity["details"] = details
    if rank == 0 and world == 1 and not args.no_extras:
        extras = run_extras(dev, peak, args.compiled_baseline)
    if extras is not None:
        extras["cfg4_long_cot_strong_scaling"] = cfg4

    if rank == 0:
        n_mb = len(mbs)
        out = {
            "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "tokens_per_step_per_gpu": tokens_per_rank, "micro_batches": n_mb,
                       "temperature": TEMPERATURE, "eps_clip": EPS_CLIP, "behav_imp_weight_cap": CAP,
                       "adv_norm": "group/group (8)", "parallelism": f"dp{world} (sequence-sharded, weak)",
                       "cache": f"inputs larger than L2: {N_LOGITS_BUFFERS} rotating logits buffers of "
                                f"{max_T}x{V} bf16 = {max_T * V * 2 / 1e9:.2f} GB each"},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "clocks": clocks,
            "fused_step": fused, "strong_scaling": cfg4, "parity": parity,
            "roofline_gae": roofline_gae(extras, peak, peak_src), "extra": extras,
            "gpu_launches": args.steps * (3 + 3 * n_mb),
            "gpu_launches_note": "per step: masked_moments+moments_finalize+masked_normalize, then per "
                                 "micro-batch logprob_fwd, ppo_loss_fwd, logprob_bwd (the scalar loss's "
                                 "backward is evaluated inside logprob_bwd)",
        }
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not all(v for k, v in parity.items() if k != "details"):
        sys.exit(3)  # a cross-GPU path disagreed with its single-GPU computation: the line says which


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200rl", choices=["b200rl", "reference"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the GAE / interval_op side metrics")
    ap.add_argument("--compiled-baseline", action="store_true",
                    help="extras: also time the torch.compile'd torch GPU path (adds Inductor compile time)")
    ap.add_argument("--no-fused", action="store_true", help="skip the fused-step extension measurement")
    ap.add_argument("--no-parity", action="store_true", help="N >= 2: skip the cross-GPU parity block")
    args = ap.parse_args()
    if args.impl == "reference":
        main_reference(args)
    else:
        main_b200rl(args)


if __name__ == "__main__":
    main()
