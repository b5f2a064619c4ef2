#!/bin/bash
# round 2, call 8 (8 GPUs): weak scaling at N=8 — ZeRO-1 (default) vs all-reduce at ViT-B/16, then ViT-L/16 and ViT-H/14
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 8 --no-cpu-baseline --no-e2e "$@" > gpurun_out/n8_$tag.json 2> gpurun_out/n8_$tag.err; echo "rc=$?" >> gpurun_out/n8_$tag.err; }
run b16_zero1 --steps 20 --warmup 5
run b16_allreduce --steps 20 --warmup 5 --zero1 0
run l16_zero1 --steps 10 --warmup 3 --model vit_l16
run h14_b128_zero1 --steps 6 --warmup 3 --model vit_h14 --batch 128
