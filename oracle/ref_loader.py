"""Loads the REAL reference functions from a reference tree: $B200RL_REFERENCE, /root/reference
(build container only) or <repo>/baseline/_ref (the `pip install --target` copy of the reference's
`areal` package, git-ignored, which travels to the GPU box with the snapshot; SURVEY §8c note 2).
TEST INFRASTRUCTURE: used by gen_golden.py to produce tests/golden/*, by
tests/test_reference_live.py when a tree is present, and by the CPU-baseline / --impl reference
legs of bench.py (which time the reference's own functions when a tree is found and fall back to
the oracle port, saying so, when none is). Nothing under -m gpu or smoke() touches this module.

No reference source is copied: modules are imported from where they lie, with stub modules for
the dependencies that are absent here (tensordict, omegaconf-based realhf.base).
"""

from __future__ import annotations

import ast
import importlib.util
import os
import sys
import types

_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _find_root() -> str:
    cands = [os.environ.get("B200RL_REFERENCE"), "/root/reference", os.path.join(_REPO, "baseline", "_ref")]
    for c in cands:
        if c and os.path.isfile(os.path.join(c, "areal", "utils", "functional.py")):
            return c
    return os.environ.get("B200RL_REFERENCE", "/root/reference")


REF_ROOT = _find_root()


def available() -> bool:
    """The reference's `areal` package (H1/H2, AdvNorm, SFT loss) can be loaded."""
    return os.path.isfile(os.path.join(REF_ROOT, "areal", "utils", "functional.py"))


def realhf_available() -> bool:
    """The legacy `realhf` tree (pygae, actor_loss_fn, interval python fallbacks) is present:
    only in a full checkout, the pip-installed package excludes it (pyproject.toml:165-168)."""
    return os.path.isfile(os.path.join(REF_ROOT, "realhf", "impl", "model", "utils", "ppo_functional.py"))


def areal_functional():
    """areal/utils/functional.py with `tensordict` stubbed and torch.compile bypassed
    (Inductor's CPU backend cannot build in this container; the eager originals are the
    'reference fp32 torch path')."""
    os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")
    if "tensordict" not in sys.modules:
        td = types.ModuleType("tensordict")
        td.TensorDict = dict
        sys.modules["tensordict"] = td
    path = os.path.join(REF_ROOT, "areal/utils/functional.py")
    spec = importlib.util.spec_from_file_location("_ref_areal_functional", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for name in ("_gather_logprobs", "_gather_logprobs_entropy"):
        fn = getattr(mod, name)
        orig = getattr(fn, "_torchdynamo_orig_callable", None)
        if orig is not None:
            setattr(mod, name, orig)
    return mod


def realhf_ppo_functional():
    """realhf/impl/model/utils/ppo_functional.py loaded by path with stub `realhf.base`."""
    for name in ("realhf", "realhf.base", "realhf.base.pkg_version"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["realhf.base.pkg_version"].is_available = lambda n: False
    sys.modules["realhf.base"].pkg_version = sys.modules["realhf.base.pkg_version"]
    sys.modules["realhf"].base = sys.modules["realhf.base"]
    path = os.path.join(REF_ROOT, "realhf/impl/model/utils/ppo_functional.py")
    spec = importlib.util.spec_from_file_location("_ref_ppo_functional", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _extract(path: str, names, extra_globals=None):
    """exec only the named top-level functions/classes of a reference file."""
    src = open(path).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body
            if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in names]
    g = dict(extra_globals or {})
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), g)
    return g


def flatten_param_py():
    """_slice_intervals_py / _set_intervals_py (realhf/impl/model/nn/flatten_param.py:65-90)."""
    import torch
    from typing import List, Tuple

    return _extract(
        os.path.join(REF_ROOT, "realhf/impl/model/nn/flatten_param.py"),
        {"_slice_intervals_py", "_set_intervals_py"},
        dict(torch=torch, List=List, Tuple=Tuple),
    )


def adv_norm_class():
    """AdvNorm (areal/engine/ppo/actor.py:370-604) bound to the real masked_normalization."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from typing import Optional

    F = areal_functional()
    g = _extract(
        os.path.join(REF_ROOT, "areal/engine/ppo/actor.py"),
        {"AdvNorm"},
        dict(torch=torch, np=np, dist=dist, Optional=Optional,
             masked_normalization=F.masked_normalization),
    )
    return g["AdvNorm"]


class _StatsRecorder:
    """Stands in for areal.utils.stats_tracker: records what the reference registers."""

    def __init__(self):
        self.denominators, self.stats = {}, {}

    def denominator(self, **kw):
        self.denominators.update(kw)

    def stat(self, denominator=None, **kw):
        for k, v in kw.items():
            self.stats[k] = (v, denominator)


def sft_loss_fn():
    """compute_packed_sft_loss (areal/engine/sft/lm_engine.py:45-83) bound to the real
    gather_logprobs and a recording stats tracker. Returns (fn, recorder)."""
    import torch

    F = areal_functional()
    rec = _StatsRecorder()
    g = _extract(
        os.path.join(REF_ROOT, "areal/engine/sft/lm_engine.py"),
        {"compute_packed_sft_loss"},
        dict(torch=torch, TensorDict=dict, gather_logprobs=F.gather_logprobs, stats_tracker=rec),
    )
    return g["compute_packed_sft_loss"], rec


def merge_intervals_ext():
    """The reference's own merge_intervals, compiled by oracle/Makefile into oracle/_ref/."""
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
    for f in os.listdir(here) if os.path.isdir(here) else []:
        if f.startswith("interval_op") and f.endswith(".so"):
            spec = importlib.util.spec_from_file_location("interval_op", os.path.join(here, f))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod.merge_intervals
    return None


def _load_ref_so(name: str):
    """A pybind module compiled by oracle/build_ref_cuda.py from the reference's own .cu file."""
    import torch  # noqa: F401  (the extension links against libtorch)

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", name + ".so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def ref_cugae_cuda():
    """csrc/cugae/gae.cu compiled for sm_100a (gae_1d_nolp_misalign, gae_2d_olp, gae_2d_nolp)."""
    return _load_ref_so("ref_cugae")


def ref_interval_op_cuda():
    """csrc/interval_op/interval_op.cu compiled for sm_100a (slice/set_intervals_{fp32,fp16,bf16})."""
    return _load_ref_so("ref_interval_op_cuda")
