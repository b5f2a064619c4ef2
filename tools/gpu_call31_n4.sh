#!/bin/bash
# 4 GPUs: bucket policy of the ZeRO-1 exchange at ViT-L/16 (1.23 GB fp32 gradient arena): 48 MB buckets (auto), two buckets
# (0.8 x arena, then the rest), ~three buckets; N=1 on the same box first
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
F="--model vit_l16 --batch 256 --steps 8 --warmup 3 --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 --no-graph --no-e2e"
timeout 300 python bench.py --gpus 1 $F > gpurun_out/bkl_n1.json 2> gpurun_out/bkl_n1.err
i=0
for mb in 0 985 500 985 0; do
  i=$((i+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 2967$i bench.py --gpus 4 $F --bucket-mb $mb > gpurun_out/bkl_n4_${mb}_$i.json 2> gpurun_out/bkl_n4_${mb}_$i.err
  echo "rc=$?" >> gpurun_out/bkl_n4_${mb}_$i.err
done
