#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for sp in 1 0; do echo "==== VITB200_ATTN_FWD_SPLIT=$sp"; VITB200_ATTN_FWD_SPLIT=$sp NEV=70 timeout 120 python tools/attn_trace_fwd.py 2>&1; done > gpurun_out/attn_trace_fwd.log
