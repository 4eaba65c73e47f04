#!/bin/bash
# round-2 GPU call B: tree accumulation sweep (levels x outputs per thread), launch lists of one G1 / G2 multiexp, ncu full of the tree kernels
mkdir -p gpurun_out
timeout 400 python tools/tree_sweep.py 22 1 > gpurun_out/r2b_sweep_g1.jsonl 2> gpurun_out/r2b_sweep_g1.err; echo "sweep g1 rc=$?"; tail -3 gpurun_out/r2b_sweep_g1.err
cat gpurun_out/r2b_sweep_g1.jsonl | cut -c1-330
timeout 400 python tools/tree_sweep.py 22 2 > gpurun_out/r2b_sweep_g2.jsonl 2> gpurun_out/r2b_sweep_g2.err; echo "sweep g2 rc=$?"; tail -3 gpurun_out/r2b_sweep_g2.err
cat gpurun_out/r2b_sweep_g2.jsonl | cut -c1-330
for g in 1 2; do
timeout 120 python tools/prof_one.py 22 0 $g > gpurun_out/r2b_plain_g$g.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2b_launches_msm_g$g.csv python tools/prof_one.py 22 0 $g > gpurun_out/r2b_ncu_g$g.log 2>&1
echo "ncu launches g$g rc=$?"
python tools/launch_summary.py gpurun_out/r2b_launches_msm_g$g.csv 100
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"tree_down0|tree_up0" -s 2 -c 4 -o gpurun_out/r2b_prof_tree python tools/prof_one.py 22 0 1 > gpurun_out/r2b_ncu2.log 2>&1
echo "ncu2 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"tree_down0|tree_up0" -s 2 -c 4 -o gpurun_out/r2b_prof_tree_g2 python tools/prof_one.py 22 0 2 > gpurun_out/r2b_ncu3.log 2>&1
echo "ncu3 rc=$?"
