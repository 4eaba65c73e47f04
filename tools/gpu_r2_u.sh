for lg in 16 18 20; do for tl in -1 0; do
echo "== 2^$lg tree_levels=$tl"
B2G16_TREE_LEVELS=$tl B2G16_MSM_TRACE=1 timeout 100 python tools/msm_trace.py $lg 1 2> /tmp/tr.err | head -3; python tools/trace_summary.py /tmp/tr.err | head -14
done; done
for lg in 16 18 20; do for tl in -1 0; do echo "== proof 2^$lg tree_levels=$tl"; B2G16_TREE_LEVELS=$tl timeout 100 python tools/prof_prove.py $lg 5 2>&1 | grep "^prove" | tail -2; done; done
