#!/bin/bash
# attention backward with dedicated drain warps (VITB200_ATTN_BWD_DRAIN bit 0 = dQ kernel, bit 1 = dK/dV kernel)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
VITB200_ATTN_BWD_DRAIN=3 timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu -k "attention or noalloc" > gpurun_out/t40.log 2>&1
echo "rc=$?" >> gpurun_out/t40.log
: > gpurun_out/attn40.log
for D in 0 1 2 3 0 3; do
  echo "== VITB200_ATTN_BWD_DRAIN=$D" >> gpurun_out/attn40.log
  VITB200_ATTN_BWD_DRAIN=$D B=256 timeout 120 python tools/attn_probe.py 2>&1 | grep -i "bwd" >> gpurun_out/attn40.log
done
