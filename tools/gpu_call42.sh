#!/bin/bash
# final state after the drain-warp attention backward: full GPU suite, smoke, A/B in the step, default bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t42.log 2>&1
echo "rc=$?" >> gpurun_out/t42.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke42.log 2>&1
B="timeout 300 python bench.py --no-cpu-baseline --no-ref-eager --no-graph --no-module-path --no-row0 --no-e2e"
VITB200_ATTN_BWD_DRAIN=0 $B > gpurun_out/bench42_nodrain.json 2> gpurun_out/bench42_a.err
$B > gpurun_out/bench42_drain.json 2> gpurun_out/bench42_b.err
timeout 400 python bench.py > gpurun_out/bench42.json 2> gpurun_out/bench42.err
echo "bench rc=$?" >> gpurun_out/bench42.err
