#!/bin/bash
# round-2 GPU call C: launch list of a 2^22 proof on the current build; ncu --set full of the in-proof tree kernels (G1 and G2)
mkdir -p gpurun_out
timeout 300 python tools/prof_prove.py 22 3 > gpurun_out/r2c_plain.log 2>&1; echo "plain rc=$?"; tail -4 gpurun_out/r2c_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2c_launches_prove_2p22.csv python tools/prof_prove.py 22 2 > gpurun_out/r2c_ncu1.log 2>&1
echo "ncu launches rc=$?"; tail -3 gpurun_out/r2c_ncu1.log
L=$(grep -o "launches [0-9]*" gpurun_out/r2c_ncu1.log | tail -1 | cut -d' ' -f2)
echo "launches per proof: $L"
python tools/launch_summary.py gpurun_out/r2c_launches_prove_2p22.csv $L | tee gpurun_out/r2c_launches_prove_2p22.txt
gzip -f gpurun_out/r2c_launches_prove_2p22.csv
timeout 900 ncu --set full --clock-control none -k regex:"tree_down0|tree_up0" -s 40 -c 4 -o /tmp/r2c_prof_tree python tools/prof_prove.py 22 1 > gpurun_out/r2c_ncu2.log 2>&1
echo "ncu2 rc=$?"; ls -la /tmp/r2c_prof_tree.ncu-rep
ncu -i /tmp/r2c_prof_tree.ncu-rep --page raw --csv > gpurun_out/r2c_prof_tree_raw.csv 2>/dev/null
[ $(stat -c %s /tmp/r2c_prof_tree.ncu-rep) -lt 30000000 ] && cp /tmp/r2c_prof_tree.ncu-rep gpurun_out/
ls -la gpurun_out
