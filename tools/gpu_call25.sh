#!/bin/bash
# ncu --set full (with source) of the fp32-output forward GEMM sites (proj, fc2: bias + dropout + residual epilogue)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-row0 --no-module-path --no-ref-eager --no-graph > gpurun_out/b25.json 2> gpurun_out/b25.err || exit 1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"float, .unsigned int.41," -s 24 -c 2 -o gpurun_out/gemm_f32epi_r02 -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-row0 --no-module-path --no-ref-eager --no-graph > gpurun_out/ncu25.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu25.log
