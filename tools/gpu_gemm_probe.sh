#!/bin/bash
# First-contact GPU script: each probe in its own process under timeout.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/gemm_probe.log
: > $L
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $L 2>&1
run() { echo "### $*" >> $L; timeout 120 python tools/gemm_probe.py "$@" >> $L 2>&1; echo "rc=$?" >> $L; }
run f32nt 300 200 100
run f32tn 300 200 1000
run nt 128 256 64
run nt 1024 768 768
run epi 1024 768 768
run nn 1024 768 768
run tn 768 768 4096
run nt 50432 768 768
run nt 50432 2304 768
run nt 50432 3072 768
run nt 50432 768 3072
run nn 50432 768 3072
run tn 768 3072 50432
run tn 2304 768 50432
run nt 8192 8192 8192
tail -n 80 $L
