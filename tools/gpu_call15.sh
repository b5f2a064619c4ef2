#!/bin/bash
# round 2, call 15: round-end rehearsal on ONE GPU — whole suite with -x, smoke(), default bench (timed), reference arm
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
python -m pytest tests -x -q -m gpu > gpurun_out/t_all15.log 2>&1
echo "all tests rc=$?" >> gpurun_out/t_all15.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke15.log 2>&1
echo "smoke rc=$?" >> gpurun_out/smoke15.log
/usr/bin/time -v python bench.py > gpurun_out/bench_v7.json 2> gpurun_out/bench_v7.err
echo "bench rc=$?" >> gpurun_out/bench_v7.err
/usr/bin/time -v python bench.py --impl reference > gpurun_out/bench_v7_ref.json 2> gpurun_out/bench_v7_ref.err
