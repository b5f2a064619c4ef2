#!/bin/bash
# epilogue read-side experiments (VITB200_GEMM_EPI_FLAGS: 1 = prefetch a tile ahead, 2 = per-lane prefetch, 4 = loads bypass L1)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
: > gpurun_out/epi26.log
for F in 2 10 8 2 10; do
  echo "== flags $F" >> gpurun_out/epi26.log
  NOREF=1 SITES="fwd proj,fwd fc2,mask+colsum" VITB200_GEMM_EPI_FLAGS=$F timeout 120 python tools/gemm_sites.py 2>&1 | grep -v "per-block\|^B=" >> gpurun_out/epi26.log
done
