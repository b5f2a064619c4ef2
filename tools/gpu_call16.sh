#!/bin/bash
# round 2, call 16: smoke() again, default bench (timed), reference arm (timed)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke16.log 2>&1
echo "smoke rc=$?" >> gpurun_out/smoke16.log
S=$SECONDS; python bench.py > gpurun_out/bench_v7.json 2> gpurun_out/bench_v7.err; echo "bench rc=$? seconds=$((SECONDS-S))" >> gpurun_out/bench_v7.err
S=$SECONDS; python bench.py --impl reference > gpurun_out/bench_v7_ref.json 2> gpurun_out/bench_v7_ref.err; echo "ref rc=$? seconds=$((SECONDS-S))" >> gpurun_out/bench_v7_ref.err
