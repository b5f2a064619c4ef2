#!/bin/bash
# round 2, call 6: whole GPU suite (parity table), bench lines of the other BASELINE configs, compute-sanitizer memcheck
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
rm -f gpurun_out/r02_parity_table.jsonl
python -m pytest tests -m gpu -q --maxfail=20 > gpurun_out/t_all6.log 2>&1
echo "all tests rc=$?" >> gpurun_out/t_all6.log
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-eager --no-module-path --no-row0"
$B --model vit_s16 --dtype f32 --batch 64 > gpurun_out/cfg_c2_s16_f32_b64.json 2> gpurun_out/cfg_c2.err
$B --model vit_l16 --batch 256 > gpurun_out/cfg_c4_l16_b256.json 2> gpurun_out/cfg_c4.err
$B --model vit_h14 --batch 128 > gpurun_out/cfg_c5a_h14_t577_b128.json 2> gpurun_out/cfg_c5a128.err
$B --model vit_h14 --batch 64 > gpurun_out/cfg_c5a_h14_t577_b64.json 2> gpurun_out/cfg_c5a64.err
$B --model vit_h14_384 --batch 64 > gpurun_out/cfg_c5b_h14_t730_b64.json 2> gpurun_out/cfg_c5b.err
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x -k "not 1000 and not 3072 and not 4099 and not 224" > gpurun_out/sanitizer_memcheck.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/sanitizer_memcheck.log
