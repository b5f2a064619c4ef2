#!/bin/bash
# round 2, call 11: attention backward, chunked (round 1) vs full-width plans, per kernel
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for f in 0 1 2 3; do echo "== VITB200_ATTN_BWD_FULL=$f"; VITB200_ATTN_BWD_FULL=$f timeout 120 python tools/attn_probe.py bwd 2>&1 | tail -2; echo "rc=$?"; done > gpurun_out/attn_bwd_full.log 2>&1
for f in 3; do for shape in "256 145 12" "256 160 12" "64 208 16" "8 224 4"; do echo "== FULL=$f $shape"; VITB200_ATTN_BWD_FULL=$f timeout 120 python tools/attn_probe.py bwd $shape 2>&1 | tail -2; done; done >> gpurun_out/attn_bwd_full.log 2>&1
