#!/bin/bash
# final-code ncu launch list of one training step (same command first without ncu), plus --set full of the weight-gradient GEMM
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-row0 --no-module-path --no-ref-eager --no-graph"
timeout 300 $CMD > gpurun_out/plain33.log 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/launches_v10.csv $CMD > gpurun_out/ncu_launches33.log 2>&1
echo "ncu launches rc=$?" >> gpurun_out/ncu_launches33.log
