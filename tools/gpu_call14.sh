#!/bin/bash
# round 2, call 14: attention with dead-warp skip — probe timings, attention tests, bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 120 python tools/attn_probe.py both > gpurun_out/attn_probe_v3.log 2>&1
timeout 400 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py tests/test_fullsize_gpu.py -m gpu -q -k "attention or noalloc or full_size" > gpurun_out/t14.log 2>&1
echo "rc=$?" >> gpurun_out/t14.log
timeout 600 python bench.py --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 --no-graph > gpurun_out/bench_v6.json 2> gpurun_out/bench_v6.err
