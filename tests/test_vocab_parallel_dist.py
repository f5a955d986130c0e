"""The vocabulary-parallel op over a real NCCL tensor-parallel group (>= 2 GPUs). On a single-GPU
box the two ranks share cuda:0 and the [T, 8] records travel over a gloo group instead."""

import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tp_check_two_ranks():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    env = dict(os.environ)
    if torch.cuda.device_count() < 2:
        env["B200RL_ONE_GPU"] = "1"
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "scripts", "tp_check.py"), "1024"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    if r.returncode != 0 and "B200RL_ONE_GPU" in env and (
            "busy or unavailable" in r.stderr or "exclusive" in r.stderr.lower()):
        pytest.skip("the GPU is in exclusive-process mode: two ranks cannot share it")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    res = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert res["ok"] and res["tp"] == 2
