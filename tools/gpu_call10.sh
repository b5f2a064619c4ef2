#!/bin/bash
# round 2, call 10: attention backward bound experiments (wrong results by construction when DBG != 0) + trace
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for d in 0 1 2 3; do echo "== VITB200_ATTN_DBG=$d"; VITB200_ATTN_DBG=$d python tools/attn_probe.py bwd 2>&1 | tail -1; done > gpurun_out/attn_bwd_dbg.log
NEV=400 python tools/attn_trace.py > gpurun_out/attn_trace_r02.log 2>&1
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 > gpurun_out/bench_v4.json 2> gpurun_out/bench_v4.err
