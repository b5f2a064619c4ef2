#!/bin/bash
# round 2, call 3: attention kernels after software-pipelining (A/B against the un-pipelined pieces), all GPU tests
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
python tools/attn_probe.py both > gpurun_out/attn_probe_v2.log 2>&1
VITB200_ATTN_DBG=4 python tools/attn_probe.py bwd > gpurun_out/attn_probe_v2_legacy.log 2>&1
NEV=200 python tools/attn_trace.py > gpurun_out/attn_trace_v2.log 2>&1
rm -f gpurun_out/r02_parity_table.jsonl
python -m pytest tests -m gpu -q --maxfail=20 > gpurun_out/t_all3.log 2>&1
echo "all tests rc=$?" >> gpurun_out/t_all3.log
python bench.py --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 > gpurun_out/bench_v2.json 2> gpurun_out/bench_v2.err
