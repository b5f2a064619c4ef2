#!/bin/bash
# ncu --set full of the final attention kernels (forward, dK/dV and dQ with drain warps), same command as the launch list
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-row0 --no-module-path --no-ref-eager --no-graph"
timeout 300 $CMD > gpurun_out/plain45.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attn_tc_fwd_kernel|attn_tc_bwd_kernel" -s 44 -c 6 -o gpurun_out/attn_final_r02 -f $CMD > gpurun_out/ncu45.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu45.log
