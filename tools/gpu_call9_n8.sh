#!/bin/bash
# round 2, call 9 (8 GPUs): calibrate the N=8 loss on ONE box — N=1 on the same box, ZeRO-1 default, one single bucket
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-ref-eager --no-module-path --no-row0 > gpurun_out/n8box_n1.json 2> gpurun_out/n8box_n1.err
run() { tag=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 8 --no-cpu-baseline --no-e2e --steps 20 --warmup 5 "$@" > gpurun_out/n8box_$tag.json 2> gpurun_out/n8box_$tag.err; echo "rc=$?" >> gpurun_out/n8box_$tag.err; }
run zero1
run zero1_onebucket --bucket-mb 100000
run allreduce_onebucket --bucket-mb 100000 --zero1 0
