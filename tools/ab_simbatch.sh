#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-sb}
for pm in 160 192 224 256; do
CM_SIMILAR_PASS_MAX=$pm BATCH_SIZES=$pm,$((2*pm)) NO_CPU=1 timeout 300 python tools/bench_similar.py > $OUT/${TAG}_bench_pm$pm.log 2>&1; echo "pass_max=$pm rc=$?"; grep "queries" $OUT/${TAG}_bench_pm$pm.log
done
