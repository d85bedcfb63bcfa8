#!/bin/bash
# Developer run: batched most_similar -- tests and timing.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-sb}
timeout 300 python -m pytest tests/test_gpu_similar.py -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; echo "similar tests rc=$?"; tail -2 $OUT/${TAG}_tests.log
CHECK_BATCH=1 NO_CPU=1 timeout 300 python tools/bench_similar.py > $OUT/${TAG}_bench.log 2>&1; echo "bench rc=$?"; grep "queries\|identical\|p50" $OUT/${TAG}_bench.log
