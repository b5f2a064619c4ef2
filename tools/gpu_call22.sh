#!/bin/bash
# round 2, call 22: ncu --set full of the attention kernels (forward with two softmax groups, both backward kernels) and
# of the GEMM call sites of one block, from the bench command
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-row0 --no-module-path --no-ref-eager --no-graph"
$CMD > gpurun_out/plain22.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attn_tc_fwd_kernel|attn_tc_bwd_kernel|attn_delta_vec_kernel" -s 12 -c 8 -o gpurun_out/attn_r02 -f $CMD > gpurun_out/ncu_attn22.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_attn22.log
