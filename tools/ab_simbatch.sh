#!/bin/bash
# Developer run: batched most_similar -- tests, timing, launch list and one full capture of the tensor-core scan.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-sb}
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; echo "gpu tests rc=$?"; tail -4 $OUT/${TAG}_tests.log
CHECK_BATCH=1 NO_CPU=1 timeout 600 python tools/bench_similar.py > $OUT/${TAG}_bench.log 2>&1; echo "bench rc=$?"; grep "queries\|identical\|p50" $OUT/${TAG}_bench.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches.csv python tools/probe_simbatch.py 128 > $OUT/${TAG}_launches.log 2>&1; echo "launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cm_simscan -s 3 -c 1 -f -o $OUT/${TAG}_simscan128 python tools/probe_simbatch.py 128 > $OUT/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
