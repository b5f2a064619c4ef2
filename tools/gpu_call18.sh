#!/bin/bash
# round 2, call 18: forward attention with two softmax groups per tile (column split) vs one
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
{ for sp in 0 1; do echo "== VITB200_ATTN_FWD_SPLIT=$sp"; VITB200_ATTN_FWD_SPLIT=$sp timeout 120 python tools/attn_probe.py fwd 2>&1 | tail -3; done
  for shape in "256 224 12" "256 225 12" "64 256 16" "256 193 12"; do echo "== SPLIT=1 $shape"; timeout 120 python tools/attn_probe.py fwd $shape 2>&1 | tail -3; done; } > gpurun_out/attn_fwd_split.log 2>&1
timeout 400 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -m gpu -q -k "attention or bf16 or fixture" > gpurun_out/t18.log 2>&1
echo "rc=$?" >> gpurun_out/t18.log
