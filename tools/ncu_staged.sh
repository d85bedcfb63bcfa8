OUT=gpurun_out
for c in 1 2; do
CM_SCORE_KERNEL=staged CM_STAGED_CFG=$c timeout 600 ncu --set full --clock-control none --import-source on -k regex:cm_score -s 40 -c 1 -f -o $OUT/st3_staged_cfg${c}_65536x1 python tools/probe_kernel.py 65536x1 reps=60 > $OUT/st3_ncu_cfg$c.log 2>&1; echo "ncu cfg$c rc=$?"
done
