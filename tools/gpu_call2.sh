#!/bin/bash
# round 2, call 2: the reference-on-GPU parity tests (all configs) + train.py unchanged, ncu --set full of LN bwd and
# AdamW (one ncu command), attention backward trace, attention probe timings
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
python -m pytest tests/test_reference_gpu.py "tests/test_trainer_gpu.py::test_adamw_range_bf16_gradients_and_fused_zero" -m gpu -q --maxfail=20 > gpurun_out/t_ref.log 2>&1
echo "ref tests rc=$?" >> gpurun_out/t_ref.log
python tools/attn_probe.py both > gpurun_out/attn_probe_base.log 2>&1
NEV=260 python tools/attn_trace.py > gpurun_out/attn_trace_base.log 2>&1
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-row0 --no-module-path --no-ref-eager"
$CMD > gpurun_out/plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ln_bwd_ring_kernel|adamw_range_kernel|ln_fwd_reg_kernel" -s 71 -c 32 -o gpurun_out/lnbwd_adamw_r02 -f $CMD > gpurun_out/ncu_ln2.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu_ln2.log
