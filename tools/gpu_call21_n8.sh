#!/bin/bash
# round 2, call 21 (8 GPUs): scaling rehearsal on ONE box with the final code — N = 1, 2, 4, 8 (driver-style launch,
# default exchange policy), then ViT-H/14 at 384x384 (T=730) on 8 GPUs
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 --no-graph > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
for n in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2955$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  echo "rc=$?" >> gpurun_out/scale_n$n.err
done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29559 bench.py --gpus 8 --steps 6 --warmup 3 --model vit_h14_384 --batch 64 --no-e2e > gpurun_out/n8_h14_384_b64.json 2> gpurun_out/n8_h14_384_b64.err
echo "rc=$?" >> gpurun_out/n8_h14_384_b64.err
