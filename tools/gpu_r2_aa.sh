#!/bin/bash
# round-2 GPU call AA: ncu --set full of every tree_up0 / tree_down0 launch (G1 and G2) inside one 2^22 proof, final build
mkdir -p gpurun_out
timeout 200 python tools/prof_prove.py 22 2 > gpurun_out/r2aa_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:"tree_down0|tree_up0" -s 46 -c 46 -o /tmp/r2aa_tree python tools/prof_prove.py 22 2 > gpurun_out/r2aa_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r2aa_ncu.log
ncu -i /tmp/r2aa_tree.ncu-rep --page raw --csv > gpurun_out/r2aa_tree_raw.csv 2>/dev/null
python tools/ncu_stage_summary.py gpurun_out/r2aa_tree_raw.csv gpurun_out/r2aa_ncu_full_tree_in_proof.json "ncu --set full --clock-control none of the 46 tree_up0 / tree_down0 launches (G1 x 36, G2 x 10) of the second of two 2^22 proofs (tools/prof_prove.py 22 2), final build"
gzip -f gpurun_out/r2aa_tree_raw.csv
