#!/bin/bash
# round 2, call 17: what does the dropout RNG cost? same box, p = 0.2 vs p = 0
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
F="--no-cpu-baseline --no-ref-eager --no-module-path --no-row0 --no-graph --no-e2e --steps 20 --warmup 5"
python bench.py $F > gpurun_out/drop_p02.json 2> gpurun_out/drop_p02.err
python bench.py $F --dropout 0.0 > gpurun_out/drop_p00.json 2> gpurun_out/drop_p00.err
python bench.py $F > gpurun_out/drop_p02b.json 2> gpurun_out/drop_p02b.err
