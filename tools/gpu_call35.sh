#!/bin/bash
# fused delta (dO GEMM epilogue): whole GPU suite, then in-step A/B against the separate delta kernel
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t35.log 2>&1
echo "rc=$?" >> gpurun_out/t35.log
B="timeout 300 python bench.py --no-cpu-baseline --no-ref-eager --no-graph --no-module-path --no-row0 --no-e2e"
VITB200_FUSED_DELTA=0 $B > gpurun_out/bench35_sep.json 2> gpurun_out/bench35_sep.err
$B > gpurun_out/bench35_fused.json 2> gpurun_out/bench35_fused.err
VITB200_FUSED_DELTA=0 $B > gpurun_out/bench35_sep2.json 2>> gpurun_out/bench35_sep.err
$B > gpurun_out/bench35_fused2.json 2>> gpurun_out/bench35_fused.err
