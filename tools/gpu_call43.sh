#!/bin/bash
# bench lines of the other BASELINE configs with the new split-K rule / epilogue defaults
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
B="timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-eager --no-module-path --no-row0 --no-graph"
$B --model vit_s16 --dtype f32 --batch 64 > gpurun_out/v12_cfg_c2_s16_f32_b64.json 2> gpurun_out/v12_cfg_c2.err
$B --model vit_l16 --batch 256 > gpurun_out/v12_cfg_c4_l16_b256.json 2> gpurun_out/v12_cfg_c4.err
$B --model vit_h14 --batch 128 > gpurun_out/v12_cfg_c5a_h14_t577_b128.json 2> gpurun_out/v12_cfg_c5a128.err
$B --model vit_h14_384 --batch 64 > gpurun_out/v12_cfg_c5b_h14_t730_b64.json 2> gpurun_out/v12_cfg_c5b.err
