#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gae_gpu.py tests/test_reference_cuda_gpu.py tests/test_zz_no_alloc.py -q -m gpu -x -p no:cacheprovider 2>&1 | tail -4
timeout 600 python scripts/time_gae_modes.py $1 2>&1 | tee gpurun_out/time_gae_modes.log
