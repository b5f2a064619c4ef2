#!/bin/bash
# 4 GPUs: gradient-bucket policy of the ZeRO-1 exchange at ViT-B/16 — one bucket (default), two buckets with the small
# one last (--bucket-mb 300 / 270), many 48 MB buckets; N=1 on the same box first
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
F="--steps 20 --warmup 5 --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 --no-graph --no-e2e"
timeout 300 python bench.py --gpus 1 $F > gpurun_out/bk_n1.json 2> gpurun_out/bk_n1.err
i=0
for mb in 0 300 270 240 48 0; do
  i=$((i+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 2966$i bench.py --gpus 4 $F --bucket-mb $mb > gpurun_out/bk_n4_${mb}_$i.json 2> gpurun_out/bk_n4_${mb}_$i.err
  echo "rc=$?" >> gpurun_out/bk_n4_${mb}_$i.err
done
