"""The C ABI from a host that has nothing to do with torch or Python: examples/cabi_host.c is
plain C (CUDA runtime only). CPU: it compiles against include/b200rl.h as C11 and links against
libb200rl.so (every entry point it names resolves). GPU: it runs GAE (serial / scan / auto),
interval slice + set and merge_intervals and checks them against scalar restatements of the
reference loops written in the C file itself."""

import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def _build(out_dir):
    exe = os.path.join(out_dir, "cabi_host")
    cmd = ["gcc", "-O2", "-std=c11", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(CUDA, "include"), os.path.join(ROOT, "examples", "cabi_host.c"),
           "-L" + os.path.join(ROOT, "areal_vllm_b200"), "-lb200rl", "-L" + os.path.join(CUDA, "lib64"),
           "-lcudart", "-lm", "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_c_host_compiles_and_links(tmp_path):
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    exe = _build(str(tmp_path))
    assert os.path.getsize(exe) > 0
    nm = subprocess.run(["nm", "-u", exe], capture_output=True, text=True).stdout
    for sym in ("b200rl_gae_1d_nolp_misalign", "b200rl_slice_intervals", "b200rl_set_intervals",
                "b200rl_merge_intervals", "b200rl_last_error"):
        assert sym in nm


@pytest.mark.gpu
def test_c_host_runs(tmp_path, cuda):
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    exe = _build(str(tmp_path))
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.pathsep.join(
        [os.path.join(ROOT, "areal_vllm_b200"), os.path.join(CUDA, "lib64"), env.get("LD_LIBRARY_PATH", "")])
    r = subprocess.run([exe], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "OK" in r.stdout and "bit-exact=1" in r.stdout
