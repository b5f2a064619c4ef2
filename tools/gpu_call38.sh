#!/bin/bash
# final state: full GPU suite, smoke, default bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t38.log 2>&1
echo "rc=$?" >> gpurun_out/t38.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke38.log 2>&1
timeout 400 python bench.py > gpurun_out/bench38.json 2> gpurun_out/bench38.err
echo "bench rc=$?" >> gpurun_out/bench38.err
