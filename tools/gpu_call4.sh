#!/bin/bash
# round 2, call 4: no-alloc entry point test alone (fresh process), then the whole GPU suite (it runs last there too),
# default bench + reference arm, ncu launch list + --set full of LayerNorm fwd/bwd and AdamW
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/t_noalloc_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/t_noalloc_alone.log
rm -f gpurun_out/r02_parity_table.jsonl
python -m pytest tests -m gpu -q --maxfail=20 > gpurun_out/t_all4.log 2>&1
echo "all tests rc=$?" >> gpurun_out/t_all4.log
python bench.py > gpurun_out/bench_v3.json 2> gpurun_out/bench_v3.err
echo "bench rc=$?" >> gpurun_out/bench_v3.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_v3_ref.json 2> gpurun_out/bench_v3_ref.err
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-row0 --no-module-path --no-ref-eager"
$CMD > gpurun_out/plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/launches_v3.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?" >> gpurun_out/ncu_launches.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ln_bwd_ring_kernel|adamw_range_kernel|ln_fwd_reg_kernel" -s 73 -c 30 -o gpurun_out/lnbwd_adamw_r02 -f $CMD > gpurun_out/ncu_ln2.log 2>&1
echo "ncu full rc=$?" >> gpurun_out/ncu_ln2.log
