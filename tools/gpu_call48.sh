#!/bin/bash
# EXACT epilogue / LayerNorm instantiations: full GPU suite, per-site A/B, smoke, default bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/t48.log 2>&1
echo "rc=$?" >> gpurun_out/t48.log
NOREF=1 VITB200_GEMM_EXACT=0 timeout 120 python tools/gemm_sites.py > gpurun_out/sites48_bounds.log 2>&1
NOREF=1 timeout 120 python tools/gemm_sites.py > gpurun_out/sites48_exact.log 2>&1
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke48.log 2>&1
timeout 400 python bench.py > gpurun_out/bench48.json 2> gpurun_out/bench48.err
echo "bench rc=$?" >> gpurun_out/bench48.err
