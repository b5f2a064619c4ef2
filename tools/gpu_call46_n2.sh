#!/bin/bash
# 2 GPUs, one box: driver-style launches with the final code — N=1 then N=2 (default exchange policy), reference arm under torchrun
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
F="--steps 20 --warmup 5 --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 --no-graph"
timeout 300 python bench.py --gpus 1 $F > gpurun_out/fin2_n1.json 2> gpurun_out/fin2_n1.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29691 bench.py --gpus 2 $F > gpurun_out/fin2_n2.json 2> gpurun_out/fin2_n2.err
echo "rc=$?" >> gpurun_out/fin2_n2.err
timeout 300 python -m pytest tests/test_trainer_gpu.py -x -q -m gpu -k "two_rank" > gpurun_out/t46.log 2>&1
echo "rc=$?" >> gpurun_out/t46.log
