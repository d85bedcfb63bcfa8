#!/bin/bash
# Developer run: blur + JPEG tests, launch list, bench.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-bl}
timeout 600 python -m pytest tests/test_gpu_blur.py tests/test_gpu_jpeg.py -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; echo "blur+jpeg tests rc=$?"; tail -3 $OUT/${TAG}_tests.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches.csv python tools/probe_jpeg.py > $OUT/${TAG}_probe.log 2>&1; echo "launches rc=$?"
timeout 300 python tools/bench_blur.py > $OUT/${TAG}_bench.log 2>&1; grep "JPEG\|page-locked" $OUT/${TAG}_bench.log
