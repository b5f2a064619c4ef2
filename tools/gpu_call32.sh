#!/bin/bash
# full GPU suite + smoke + both bench arms with the final code of the round (driver-style defaults)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
S=$SECONDS
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t32.log 2>&1
echo "rc=$? secs=$((SECONDS-S))" >> gpurun_out/t32.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke32.log 2>&1
S=$SECONDS
timeout 400 python bench.py --impl reference > gpurun_out/bench32_ref.json 2> gpurun_out/bench32_ref.err
echo "ref rc=$? secs=$((SECONDS-S))" >> gpurun_out/bench32_ref.err
S=$SECONDS
timeout 400 python bench.py > gpurun_out/bench32.json 2> gpurun_out/bench32.err
echo "bench rc=$? secs=$((SECONDS-S))" >> gpurun_out/bench32.err
