"""Compiles the reference's two CUDA extensions for sm_100a FROM WHERE THEY LIE under
/root/reference into oracle/_ref/ (git-ignored, travels to the GPU box). TEST INFRASTRUCTURE.

They are the GPU-side "kernels to beat" and a second parity pin on the box: tests/
test_reference_cuda_gpu.py compares our kernels with the real reference kernels, and
scripts/time_misc.py times both. Only runs where the reference tree exists; no source is copied.
"""

import os
import sys

import torch
from torch.utils import cpp_extension

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B200RL_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")


def main():
    if not os.path.isdir(os.path.join(REF, "csrc")):
        print("reference tree not present: nothing to build")
        return 0
    os.makedirs(OUT, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    flags = ["-O3", "-gencode", "arch=compute_100a,code=sm_100a"]
    for name, src in (("ref_cugae", "csrc/cugae/gae.cu"),
                      ("ref_interval_op_cuda", "csrc/interval_op/interval_op.cu")):
        target = os.path.join(OUT, name + ".so")
        if os.path.exists(target) and os.path.getmtime(target) > os.path.getmtime(os.path.join(REF, src)):
            print("up to date:", target)
            continue
        bdir = os.path.join("/tmp", "b200rl_refbuild_" + name)
        os.makedirs(bdir, exist_ok=True)
        cpp_extension.load(name=name, sources=[os.path.join(REF, src)], extra_cuda_cflags=flags,
                           build_directory=bdir, is_python_module=False, verbose=False)
        import shutil
        shutil.copy(os.path.join(bdir, name + ".so"), target)
        print("built", target)
    return 0


if __name__ == "__main__":
    sys.exit(main())
