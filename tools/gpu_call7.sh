#!/bin/bash
# round 2, call 7: forward attention A/B (lock-step vs skewed issue order), attention + trainer tests
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
VITB200_ATTN_FWD_ORDER=0 python tools/attn_probe.py fwd > gpurun_out/attn_fwd_order0.log 2>&1
VITB200_ATTN_FWD_ORDER=1 python tools/attn_probe.py fwd > gpurun_out/attn_fwd_order1.log 2>&1
VITB200_ATTN_FWD_ORDER=1 python tools/attn_probe.py fwd 256 129 12 >> gpurun_out/attn_fwd_order1.log 2>&1
VITB200_ATTN_FWD_ORDER=1 python tools/attn_probe.py fwd 64 100 16 >> gpurun_out/attn_fwd_order1.log 2>&1
python -m pytest tests/test_kernels_gpu.py tests/test_trainer_gpu.py tests/test_zz_no_alloc.py -m gpu -q -k "attention or zero1 or noalloc or two_rank" > gpurun_out/t7.log 2>&1
echo "rc=$?" >> gpurun_out/t7.log
