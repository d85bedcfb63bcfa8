#!/bin/bash
# Developer A/B: planar against interleaved box blur (CM_BLUR_PLANAR=0): tests, kernel times from an ncu launch list.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-bl}
timeout 600 python -m pytest tests/test_gpu_blur.py tests/test_gpu_jpeg.py -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; echo "blur+jpeg tests rc=$?"; tail -3 $OUT/${TAG}_tests.log
for m in 1 0; do
CM_BLUR_PLANAR=$m timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches_p$m.csv python tools/probe_jpeg.py > $OUT/${TAG}_probe_p$m.log 2>&1; echo "launches planar=$m rc=$?"
done
timeout 300 python tools/bench_blur.py > $OUT/${TAG}_bench.log 2>&1; grep "JPEG\|page-locked" $OUT/${TAG}_bench.log
