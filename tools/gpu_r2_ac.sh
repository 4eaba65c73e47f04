#!/bin/bash
# round-2 GPU call AC: G2 addition pass capped to one CTA per SM (dummy dynamic shared memory) so that G1 kernels co-reside
for kb in 0 120 80; do
echo "== B2G16_G2_TREE_SMEM_KB=$kb"
B2G16_G2_TREE_SMEM_KB=$kb timeout 200 python tools/prof_prove.py 22 7 2>&1 | grep "^prove" | tail -4
done
