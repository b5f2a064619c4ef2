#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
{ for sp in 1 0; do echo "== lean issue, VITB200_ATTN_FWD_SPLIT=$sp"; VITB200_ATTN_FWD_SPLIT=$sp timeout 120 python tools/attn_probe.py fwd 2>&1 | tail -3; done
  for shape in "256 224 12" "64 256 16" "256 129 12" "32 100 4"; do echo "== $shape"; timeout 120 python tools/attn_probe.py fwd $shape 2>&1 | tail -3; done; } > gpurun_out/attn_fwd_lean.log 2>&1
NEV=10 timeout 120 python tools/attn_trace_fwd.py > gpurun_out/attn_trace_fwd2.log 2>&1
timeout 400 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -m gpu -q -k "attention or bf16 or fixture" > gpurun_out/t20.log 2>&1
echo "rc=$?" >> gpurun_out/t20.log
