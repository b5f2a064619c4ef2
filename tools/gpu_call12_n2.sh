#!/bin/bash
# round 2, call 12 (2 GPUs): whole GPU suite (ZeRO-1 test over NCCL), smoke(), N=2 bench with the auto bucket policy
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
python -m pytest tests -m gpu -q -x > gpurun_out/t_all12.log 2>&1
echo "all tests rc=$?" >> gpurun_out/t_all12.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke12.log 2>&1
echo "smoke rc=$?" >> gpurun_out/smoke12.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/n2_auto.json 2> gpurun_out/n2_auto.err
echo "rc=$?" >> gpurun_out/n2_auto.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/n2_ref.json 2> gpurun_out/n2_ref.err
echo "rc=$?" >> gpurun_out/n2_ref.err
