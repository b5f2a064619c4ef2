#!/bin/bash
# 8 GPUs, one box: N=1 then N=8 with the final code and default exchange policy (two buckets), driver-style launch
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
F="--steps 20 --warmup 5 --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 --no-graph"
timeout 300 python bench.py --gpus 1 $F > gpurun_out/fin_n1.json 2> gpurun_out/fin_n1.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29681 bench.py --gpus 8 $F > gpurun_out/fin_n8.json 2> gpurun_out/fin_n8.err
echo "rc=$?" >> gpurun_out/fin_n8.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29682 bench.py --gpus 8 $F --bucket-mb 100000 > gpurun_out/fin_n8_onebucket.json 2> gpurun_out/fin_n8_onebucket.err
echo "rc=$?" >> gpurun_out/fin_n8_onebucket.err
