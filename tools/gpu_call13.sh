#!/bin/bash
# round 2, call 13: CUDA-graph capture of the step — test, then bench with the graph side number
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_trainer_gpu.py -m gpu -q -k "graph" > gpurun_out/t_graph.log 2>&1
echo "rc=$?" >> gpurun_out/t_graph.log
timeout 300 python -m pytest tests/test_trainer_gpu.py tests/test_model_gpu.py -m gpu -q > gpurun_out/t13.log 2>&1
echo "rc=$?" >> gpurun_out/t13.log
timeout 600 python bench.py --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 > gpurun_out/bench_v5.json 2> gpurun_out/bench_v5.err
echo "rc=$?" >> gpurun_out/bench_v5.err
