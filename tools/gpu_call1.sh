#!/bin/bash
# round 2, call 1: GPU tests (incl. the new reference-on-GPU parity + train.py tests), default bench, reference arm,
# ncu --set full of the bandwidth kernels (LayerNorm fwd/bwd, AdamW)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
python -m pytest tests -m gpu -q --maxfail=12 -x -k "reference_gpu or no_grad or attention_probs_export or second_backward or bad_labels or adamw_range" > gpurun_out/t_new.log 2>&1
echo "new tests rc=$?" >> gpurun_out/t_new.log
python -m pytest tests -m gpu -q --maxfail=12 --deselect tests/test_reference_gpu.py > gpurun_out/t_all.log 2>&1
echo "all tests rc=$?" >> gpurun_out/t_all.log
python bench.py > gpurun_out/bench_v1.json 2> gpurun_out/bench_v1.err
echo "bench rc=$?" >> gpurun_out/bench_v1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_v1_ref.json 2> gpurun_out/bench_v1_ref.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ln_fwd_reg_kernel|ln_bwd_ring_kernel|adamw_range_kernel" -s 60 -c 7 -o gpurun_out/ln_adamw_r02 -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-row0 --no-module-path --no-ref-eager > gpurun_out/ncu_ln.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_ln.log
