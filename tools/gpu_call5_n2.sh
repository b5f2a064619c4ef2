#!/bin/bash
# round 2, call 5 (2 GPUs): ZeRO-1 over NCCL — two-rank test, then bench at N=2 with ZeRO-1 (default), all-reduce, CTA cap
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/n2_gpus.txt
python -m pytest tests/test_trainer_gpu.py -m gpu -q -k "zero1 or two_rank" > gpurun_out/t_n2.log 2>&1
echo "rc=$?" >> gpurun_out/t_n2.log
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline "$@" > gpurun_out/n2_$tag.json 2> gpurun_out/n2_$tag.err; echo "rc=$?" >> gpurun_out/n2_$tag.err; }
run zero1
run allreduce --zero1 0
run zero1_cta8 --nccl-max-ctas 8
run zero1_fp32comm --grad-comm fp32
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 > gpurun_out/n2_single.json 2> gpurun_out/n2_single.err
