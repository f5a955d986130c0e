"""No-GPU checks: the C-ABI library loads and exports every symbol include/b200rl.h declares,
argument validation returns error codes without touching a device, host-side logic
(merge_intervals, sharding, fused all-reduce over gloo with world_size 2), and the drop-in
modules keep the reference's signatures and refuse CPU tensors loudly."""

import inspect
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200rl.h")


def _declared():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b200rl_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported_and_bound():
    from areal_vllm_b200 import _lib

    names = _declared()
    assert len(names) >= 17
    nm = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\sT\s+(b200rl_[a-z0-9_]+)", nm))
    assert set(names) <= exported, set(names) - exported
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert _lib.lib.b200rl_version() >= 100
    # only our own entry points are exported with the b200rl_ prefix, all of them declared
    assert exported == set(names), exported - set(names)


def test_argument_errors_without_a_device():
    from areal_vllm_b200 import _lib

    L = _lib.lib
    assert L.b200rl_logprob_entropy_fwd(None, 0, 8, None, -1, 8, 1.0, None, None, None, None, None, None) == -1
    assert "bad shape" in _lib.last_error()
    assert L.b200rl_logprob_entropy_fwd(None, 0, 8, None, 4, 8, 0.0, None, None, None, None, None, None) == -1
    assert "temperature" in _lib.last_error()
    assert L.b200rl_logprob_entropy_fwd(None, 0, 8, None, 0, 8, 1.0, None, None, None, None, None, None) == 0  # T=0
    assert L.b200rl_gae_1d_nolp_misalign(None, None, None, None, None, None, 3, 10, 1.0, 1.0, 0, None, 0, None) == -1
    assert L.b200rl_slice_intervals(None, 10, None, 5, None, 3, 3, None, 0, 0, None) == -1  # elem_size 3
    assert L.b200rl_masked_moments(None, None, 10, 3, None, 1, None, None, 0, None) == -1   # 10 % 3
    assert L.b200rl_interval_workspace_bytes(100000, 1 << 30, 2) > 100000 * 8
    assert L.b200rl_gae_workspace_bytes(4, 1 << 20) >= (1 << 20) // 1024 * 8
    assert L.b200rl_ppo_loss_workspace_bytes(10) >= 256 + 16
    # entry points added after the first milestone: same convention (code + message, no device touched)
    assert L.b200rl_critic_loss_fwd(None, None, None, None, 5, 0.2, 7, 10.0, None, None, None, None, None, None, 0, None) == -1
    assert L.b200rl_critic_loss_workspace_bytes(1 << 20) >= 256 + 16
    assert L.b200rl_packed_rewards(None, None, None, None, None, 2, 5, 0.1, 1.0, 0, None, None, None) == -1
    assert L.b200rl_packed_rewards(None, None, None, None, None, 2, 0, 0.1, 1.0, 0, None, None, None) == 0  # T=0
    assert L.b200rl_sft_loss_fwd(None, None, None, 0, 5, 2, None, None, None, None, None, None, None, 0, None) == -1
    assert L.b200rl_sft_loss_workspace_bytes(1 << 20, 100) >= 256 + 16
    assert L.b200rl_logprob_shard_fwd(None, 0, 8, None, 4, 8, 1.0, 1, 1, None, None) == -1
    assert "partial" in _lib.last_error()
    assert L.b200rl_logprob_shard_combine(None, 0, 4, 1.0, None, None, None, None, None, None) == -1
    assert L.b200rl_move_intervals(None, 10, None, 10, None, 1, None, 1, 5, 3, None, 0, 0, None) == -1  # elem_size 3
    assert L.b200rl_move_intervals(None, 10, None, 10, None, 0, None, 0, 0, 2, None, 0, 0, None) == 0   # nothing to move
    assert L.b200rl_move_workspace_bytes(1000, 10, 1 << 20, 2) > 1010 * 8
    assert L.b200rl_gae_2d_olp(None, None, None, None, None, None, None, None, 4, 10, 0, 1.0, 1.0, None) == -1
    assert L.b200rl_gae_2d_olp(None, None, None, None, None, None, None, None, 0, 10, 0, 1.0, 1.0, None) == 0
    assert L.b200rl_ipc_export(None, None, None) == -1


def test_merge_intervals_host():
    from areal_vllm_b200.flatten_param import merge_intervals
    from conftest import load_golden

    g = load_golden("h4_intervals")
    for a, b in (("merge_in", "merge_out"), ("merge_in2", "merge_out2")):
        assert merge_intervals([tuple(x) for x in g[a].tolist()]) == [tuple(x) for x in g[b].tolist()]
    assert merge_intervals([]) == []
    assert merge_intervals([(3, 7)]) == [(3, 7)]
    assert merge_intervals([(0, 1), (1, 2), (2, 3)]) == [(0, 3)]
    assert merge_intervals([(5, 6), (0, 5)]) == [(5, 6), (0, 5)]  # order preserving, no sort


def test_shard_groups_by_tokens():
    from areal_vllm_b200.dist import shard_groups_by_tokens

    rng = np.random.default_rng(0)
    lens = rng.integers(1024, 16384, 120).tolist()
    for world in (1, 2, 4, 8):
        parts = shard_groups_by_tokens(lens, 8, world)
        assert sorted(i for p in parts for i in p) == list(range(120))
        for p in parts:  # whole groups of adjacent sequences
            assert len(p) % 8 == 0
            for k in range(0, len(p), 8):
                assert p[k] % 8 == 0 and p[k:k + 8] == list(range(p[k], p[k] + 8))
        tok = [sum(lens[i] for i in p) for p in parts]
        assert max(tok) <= 1.0 * sum(tok) / world + max(sum(lens[g * 8:(g + 1) * 8]) for g in range(15))
    with pytest.raises(ValueError):
        shard_groups_by_tokens(lens[:7], 8, 2)


def test_shard_sequences_by_tokens():
    """Sequence-granular, token-balanced split (reference: allocate_balanced_mbs, data.py:219-243)."""
    from areal_vllm_b200.dist import shard_sequences_by_tokens

    rng = np.random.default_rng(4)
    lens = []
    while sum(lens) < 1_048_576:  # BASELINE config 4
        lens.append(int(rng.integers(1024, 16385)))
    lens[-1] -= sum(lens) - 1_048_576
    for world in (1, 2, 4, 8):
        parts = shard_sequences_by_tokens(lens, world)
        assert sorted(i for p in parts for i in p) == list(range(len(lens)))
        tok = [sum(lens[i] for i in p) for p in parts]
        assert max(tok) <= 1.01 * sum(tok) / world, (world, tok)
        assert parts == shard_sequences_by_tokens(lens, world)  # deterministic
    # ragged corners: fewer sequences than ranks, empty sequences, one giant sequence
    assert shard_sequences_by_tokens([5], 4) == [[0], [], [], []]
    parts = shard_sequences_by_tokens([0, 0, 7, 0], 2)
    assert sorted(i for p in parts for i in p) == [0, 1, 2, 3]
    parts = shard_sequences_by_tokens([100, 1, 1, 1], 2)
    assert sorted(map(sorted, parts)) == [[0], [1, 2, 3]]
    with pytest.raises(ValueError):
        shard_sequences_by_tokens([1, 2], 0)


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from areal_vllm_b200.dist import allreduce_stats, shard_groups_by_tokens

    lens = list(range(10, 10 + 32))
    mine = shard_groups_by_tokens(lens, 4, world)[rank]
    mom = torch.tensor([[float(len(mine)), float(sum(lens[i] for i in mine)), 1.5 * (rank + 1)]], dtype=torch.float64)
    w = torch.tensor([float(rank + 1)], dtype=torch.float32)
    mom_g, w_g = allreduce_stats(mom, w)
    q.put((rank, mom_g.tolist(), w_g.tolist(), w_g.dtype == torch.float32))
    dist.destroy_process_group()


def test_allreduce_stats_gloo_world2():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, mom, w, keeps_dtype in res:
        assert mom == [[32.0, float(sum(range(10, 42))), 4.5]]
        assert w == [3.0] and keeps_dtype


EXPECTED = {
    "functional": {
        "gather_logprobs": ["logits", "labels", "temperature", "chunk_size"],
        "gather_logprobs_entropy": ["logits", "labels", "temperature", "chunk_size"],
        "masked_normalization": ["x", "mask", "dim", "unbiased", "eps", "high_precision", "all_reduce", "reduce_group"],
        "ppo_actor_loss_fn": ["logprobs", "proximal_logprobs", "old_logprobs", "advantages", "eps_clip",
                              "loss_mask", "eps_clip_higher", "c_clip", "behav_imp_weight_cap"],
        "dynamic_sampling": ["data", "group_size"],
        "reward_overlong_penalty": ["data", "overlong_tokens", "overlong_penalty_factor", "max_response_length"],
    },
    "ppo_functional": {
        "cugae1d_nolp_misalign_func": ["rewards", "values", "cu_seqlens", "truncate", "gamma", "lam"],
        "pygae1d_nolp_misalign": ["rewards", "values", "cu_seqlens_", "bootstrap", "gamma", "lam"],
        "get_packed_advantages_and_returns": ["gamma", "lam", "values", "rewards", "short1cu_seqlens", "seq_no_eos_mask"],
        "actor_loss_fn": ["logprobs", "old_logprobs", "advantages", "eps_clip", "loss_mask", "c_clip",
                          "proximal_logprobs", "behav_imp_weight_cap"],
    },
    "flatten_param": {
        "slice_intervals": ["src", "intervals_cpu", "intervals_cuda", "output_size", "max_interval_size"],
        "set_intervals": ["src", "dst", "intervals_cpu", "intervals_cuda", "max_interval_size"],
    },
}


def test_dropin_signatures():
    import areal_vllm_b200 as pkg

    for modname, fns in EXPECTED.items():
        mod = getattr(pkg, modname)
        for fn, params in fns.items():
            got = list(inspect.signature(getattr(mod, fn)).parameters)
            assert got[: len(params)] == params, (modname, fn, got)
    sig = inspect.signature(pkg.functional.gather_logprobs)
    assert sig.parameters["temperature"].default == 1.0 and sig.parameters["chunk_size"].default == 1024
    sig = inspect.signature(pkg.functional.masked_normalization)
    assert sig.parameters["eps"].default == 1e-5 and sig.parameters["high_precision"].default is True
    assert list(inspect.signature(pkg.actor.grpo_loss_fn).parameters)[:7] == [
        "logits", "input_data", "temperature", "eps_clip", "eps_clip_higher", "c_clip", "behav_imp_weight_cap"]
    assert list(inspect.signature(pkg.actor.AdvNorm.__call__).parameters) == [
        "self", "advantages", "loss_mask", "eps", "unbiased", "high_precision", "reduce_group"]


def test_signatures_match_live_reference():
    """When the reference tree is present (build container), compare against the real thing."""
    from oracle import ref_loader as R

    if not (R.available() and R.realhf_available()):
        pytest.skip("full reference tree not present")
    import areal_vllm_b200 as pkg

    F, P = R.areal_functional(), R.realhf_ppo_functional()
    for name in EXPECTED["functional"]:
        ref = inspect.signature(getattr(F, name))
        got = inspect.signature(getattr(pkg.functional, name))
        assert list(ref.parameters) == list(got.parameters)[: len(ref.parameters)], name
        for k, p in ref.parameters.items():
            assert got.parameters[k].default == p.default, (name, k)
    for name in EXPECTED["ppo_functional"]:
        ref = inspect.signature(getattr(P, name))
        got = inspect.signature(getattr(pkg.ppo_functional, name))
        assert list(ref.parameters) == list(got.parameters)[: len(ref.parameters)], name


def test_no_cpu_fallback():
    import areal_vllm_b200 as pkg

    x = torch.randn(4, 16).bfloat16()
    y = torch.zeros(4, dtype=torch.long)
    f = torch.zeros(4)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.functional.gather_logprobs(x, y)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.functional.gather_logprobs_entropy(x, y)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.functional.masked_normalization(f)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.functional.ppo_actor_loss_fn(f, f, f, f, 0.2, f.bool())
    with pytest.raises(RuntimeError, match="CPU"):
        pkg.ppo_functional.pygae1d_nolp_misalign(f, torch.zeros(5), torch.tensor([0, 4]), torch.ones(1), 1.0, 1.0)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.flatten_param.slice_intervals(f, [(0, 2)])
    # the product never imports the oracle
    import ast
    for root, _, files in os.walk(os.path.join(ROOT, "areal_vllm_b200")):
        for fn in files:
            if fn.endswith(".py"):
                tree = ast.parse(open(os.path.join(root, fn)).read())
                for node in ast.walk(tree):
                    mods = []
                    if isinstance(node, ast.Import):
                        mods = [a.name for a in node.names]
                    elif isinstance(node, ast.ImportFrom) and node.module:
                        mods = [node.module]
                    assert not any(m.split(".")[0] == "oracle" for m in mods), (fn, mods)


def test_plumbing_functions_cpu():
    """dynamic_sampling / reward_overlong_penalty are tensor plumbing, not kernels (functional.py:191-259)."""
    import areal_vllm_b200 as pkg

    class TD(dict):
        def __getitem__(self, k):
            if isinstance(k, torch.Tensor):
                return TD({n: v[k] for n, v in self.items()})
            return dict.__getitem__(self, k)

    data = TD(rewards=torch.tensor([1., 1., 0., 1., 2., 2., 3., 3.]), x=torch.arange(8))
    out, stats = pkg.functional.dynamic_sampling(data, 2)
    assert stats == dict(n_group_keeped=1, n_group_filtered=3) and out["x"].tolist() == [2, 3]
    same = TD(rewards=torch.ones(4), x=torch.arange(4))
    assert pkg.functional.dynamic_sampling(same, 2) is same          # early exit returns data alone
    with pytest.warns(UserWarning):
        assert pkg.functional.dynamic_sampling(same, 3) is same
    d = dict(rewards=torch.tensor([1.0, 1.0]), input_ids=torch.zeros(2, 10),
             loss_mask=torch.tensor([[1.] * 9 + [0.], [1.] * 3 + [0.] * 7]))
    out = pkg.functional.reward_overlong_penalty(d, overlong_tokens=4, overlong_penalty_factor=1.0, max_response_length=10)
    np.testing.assert_allclose(out["rewards"].numpy(), [1.0 - 3 / 4, 1.0])
    from oracle import ref_loader as R
    if R.available():
        F = R.areal_functional()
        d2 = dict(rewards=torch.tensor([1.0, 1.0]), input_ids=torch.zeros(2, 10), loss_mask=d["loss_mask"])
        ref = F.reward_overlong_penalty(d2, overlong_tokens=4, overlong_penalty_factor=1.0, max_response_length=10)
        np.testing.assert_allclose(out["rewards"].numpy(), ref["rewards"].numpy())


def test_new_dropin_signatures_match_live_reference():
    """critic_loss_fn / get_packed_rewards / compute_packed_sft_loss / Megatron grpo_loss_fn."""
    from oracle import ref_loader as R

    import areal_vllm_b200 as pkg

    want = {
        "critic_loss_fn": ["value", "old_value", "target_value", "value_eps_clip", "loss_mask", "loss_fn_type"],
        "get_packed_rewards": ["kl_ctl", "clip_reward_value", "log_probs", "ref_log_probs", "reward_score",
                               "short1cu_seqlens", "seq_no_eos_mask", "mask_no_eos_with_zero"],
    }
    for name, params in want.items():
        assert list(inspect.signature(getattr(pkg.ppo_functional, name)).parameters) == params
    assert list(inspect.signature(pkg.lm_engine.compute_packed_sft_loss).parameters)[:2] == ["logits", "input_"]
    assert list(inspect.signature(pkg.mcore_functional.grpo_loss_fn).parameters)[:6] == [
        "logits", "input_data", "temperature", "eps_clip", "c_clip", "behav_imp_weight_cap"]
    if not R.realhf_available():
        return
    P = R.realhf_ppo_functional()
    for name, params in want.items():
        ref = inspect.signature(getattr(P, name))
        got = inspect.signature(getattr(pkg.ppo_functional, name))
        assert list(ref.parameters) == params
        for k, p in ref.parameters.items():
            assert got.parameters[k].default == p.default, (name, k)


def _np_partial(x, labels, start, scale):
    """numpy statement of one shard's 8-float record (M, S, W, x_y, max, min, 0, 0)."""
    z = x.astype(np.float64) * scale
    M = z.max(-1)
    e = np.exp2(z - M[:, None])
    S = e.sum(-1)
    W = (e * (z - M[:, None])).sum(-1)
    loc = labels - start
    own = (loc >= 0) & (loc < x.shape[1])
    xy = np.where(own, x[np.arange(len(x)), np.clip(loc, 0, x.shape[1] - 1)], np.nan)
    return np.stack([M, S, W, xy, x.max(-1), x.min(-1), 0 * M, 0 * M], -1).astype(np.float32)


def _np_combine(parts, scale):
    parts = parts.astype(np.float64)
    M = parts[:, :, 0].max(0)
    f = np.exp2(parts[:, :, 0] - M)
    S = (parts[:, :, 1] * f).sum(0)
    W = (f * ((parts[:, :, 0] - M) * parts[:, :, 1] + parts[:, :, 2])).sum(0)
    xy = np.nansum(parts[:, :, 3], 0)
    ln2 = np.log(2.0)
    return ((xy * scale - M) - np.log2(S)) * ln2, ln2 * (np.log2(S) - W / S)


def _tp_worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from areal_vllm_b200.mcore_functional import gather_partials, vocab_range

    rng = np.random.default_rng(5)  # same on every rank
    T, V, temp = 33, 96, 0.7
    x = (rng.standard_normal((T, V)) * 3).astype(np.float32)
    labels = rng.integers(0, V, T)
    s, e = vocab_range(V // world, rank, world)
    part = torch.from_numpy(_np_partial(x[:, s:e], labels, s, np.log2(np.e) / temp))
    allp = gather_partials(part, dist.group.WORLD)
    assert allp.shape == (world, T, 8)
    assert torch.equal(allp[rank].view(torch.int32), part.view(torch.int32))  # bits: x_y is nan off-shard
    lp, ent = _np_combine(allp.numpy(), np.log2(np.e) / temp)
    q.put((rank, (s, e), lp.tolist(), ent.tolist()))
    dist.destroy_process_group()


def test_vocab_parallel_exchange_gloo_world2():
    """The N>1 host path of the vocabulary-parallel op: shard ranges, label ownership and the one
    all-gather, with the record algebra restated in numpy and checked against the oracle."""
    import torch.multiprocessing as mp

    from areal_vllm_b200.mcore_functional import gather_partials, vocab_range
    from oracle import torch_ref

    assert vocab_range(100, 0, 4) == (0, 100) and vocab_range(100, 3, 4) == (300, 400)
    with pytest.raises(ValueError):
        vocab_range(100, 4, 4)
    one = torch.zeros(5, 8)
    assert gather_partials(one, None).shape == (1, 5, 8)

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_tp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert res[0][1] == (0, 48) and res[1][1] == (48, 96)
    rng = np.random.default_rng(5)
    x = (rng.standard_normal((33, 96)) * 3).astype(np.float32)
    labels = rng.integers(0, 96, 33)
    lp64, ent64 = torch_ref.logprobs_entropy_f64(torch.from_numpy(x), torch.from_numpy(labels), 0.7)
    for _, _, lp, ent in res:  # both ranks hold the same full-vocabulary answer
        np.testing.assert_allclose(lp, lp64.numpy(), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(ent, ent64.numpy(), rtol=1e-5, atol=1e-6)


def test_gae_mode_codes_match_the_header():
    """ppo_functional's mode names carry the integer codes include/b200rl.h defines (the C side
    switches on them), and every forced SERIAL kernel of the header has a name."""
    import re

    from areal_vllm_b200 import _lib, ppo_functional as PF

    hdr = open(os.path.join(ROOT, "include", "b200rl.h")).read()
    codes = {m.group(1).lower(): int(m.group(2)) for m in re.finditer(r"#define\s+B200RL_GAE_(\w+)\s+(\d+)", hdr)}
    assert codes["auto"] == _lib.GAE_AUTO and codes["serial"] == _lib.GAE_SERIAL and codes["scan"] == _lib.GAE_SCAN
    assert len(set(codes.values())) == len(codes)
    for name, code in codes.items():
        assert PF._MODES[name] == code, name
    assert set(PF._MODES) == set(codes)


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference tree is only present in the build container")
def test_reference_citations_point_at_existing_lines():
    """Every `path:line[-line]` citation of a reference file in the header, the docs, the package
    and the oracle names a file that exists in the reference tree and a range inside it."""
    pat = re.compile(r"((?:areal|realhf|csrc)/[\w/.\-]+\.(?:py|cu|cpp|h|cuh)):(\d+)(?:-(\d+))?")
    files = ["include/b200rl.h", "DESIGN.md", "INTEGRATION.md", "BASELINE.md", "README.md"]
    for sub, exts in (("areal_vllm_b200", (".py",)), ("areal_vllm_b200/csrc", (".cu", ".cuh", ".cpp")),
                      ("oracle", (".py", ".c"))):
        files += [os.path.join(sub, f) for f in sorted(os.listdir(os.path.join(ROOT, sub))) if f.endswith(exts)]
    bad, n = [], 0
    for fn in files:
        for m in pat.finditer(open(os.path.join(ROOT, fn)).read()):
            path, a, b = m.group(1), int(m.group(2)), int(m.group(3) or m.group(2))
            ref = os.path.join("/root/reference", path)
            if not os.path.exists(ref):
                if path.startswith("csrc/") and os.path.exists(os.path.join(ROOT, "areal_vllm_b200", path)):
                    continue  # one of this repo's own kernels
                bad.append((fn, m.group(0), "no such file"))
                continue
            n += 1
            with open(ref, errors="ignore") as f:
                length = sum(1 for _ in f)
            if a > b or b > length:
                bad.append((fn, m.group(0), f"file has {length} lines"))
    assert n > 80 and not bad, bad


def test_merge_intervals_properties():
    """Size-independent properties of merge_intervals against a plain-python statement of the
    reference loop (interval_op.cpp:4-25): neighbours that touch are joined in the order given,
    the covered elements and their order do not change, and merging is idempotent."""
    from hypothesis import given, settings, strategies as st

    from areal_vllm_b200.flatten_param import merge_intervals

    def model(iv):
        out = []
        for a, b in iv:
            if out and out[-1][1] == a:
                out[-1] = (out[-1][0], b)
            else:
                out.append((a, b))
        return out

    pairs = st.lists(st.tuples(st.integers(0, 40), st.integers(0, 12)).map(lambda t: (t[0], t[0] + t[1])), max_size=60)
    from oracle import ref_loader

    ref_merge = ref_loader.merge_intervals_ext()  # the reference's own interval_op.cpp, when it was built here

    @settings(max_examples=300, deadline=None)
    @given(pairs)
    def check(iv):
        got = merge_intervals(iv)
        assert got == model(iv)
        if ref_merge is not None:
            assert got == [tuple(x) for x in ref_merge(iv)]
        assert merge_intervals(got) == got
        flat = lambda L: [x for a, b in L for x in range(a, b)]
        assert flat(got) == flat(iv)

    check()
    # a chain of a million touching intervals collapses into one
    n = 1_000_000
    big = [(i, i + 1) for i in range(n)]
    assert merge_intervals(big) == [(0, n)]
