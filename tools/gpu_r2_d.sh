#!/bin/bash
# round-2 GPU call D: split / Karatsuba / lazy-reduction multiplier microbench; in-proof sweep of the tree knobs
mkdir -p gpurun_out
timeout 120 tools/mul_bench2 | tee gpurun_out/r2d_mul_bench2.jsonl
for cfg in "-1 16 32" "-1 16 16" "-1 16 8" "-1 16 4" "-1 32 8" "3 16 8" "4 16 8" "4 32 8" "3 32 32" "4 16 32"; do
set -- $cfg
echo "== tree_levels=$1 tree_k=$2 tree_k2=$3"
B2G16_TREE_LEVELS=$1 B2G16_TREE_K=$2 B2G16_TREE_K2=$3 timeout 200 python tools/prof_prove.py 22 4 2>&1 | grep "^prove" | tail -2
done | tee gpurun_out/r2d_tree_knobs_in_proof.txt
