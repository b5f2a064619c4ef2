#!/bin/bash
# full GPU suite with the new split-K rule + epilogue defaults, smoke, default bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
S=$SECONDS
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t27.log 2>&1
echo "rc=$? secs=$((SECONDS-S))" >> gpurun_out/t27.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke27.log 2>&1
S=$SECONDS
timeout 400 python bench.py > gpurun_out/bench27.json 2> gpurun_out/bench27.err
echo "bench rc=$? secs=$((SECONDS-S))" >> gpurun_out/bench27.err
