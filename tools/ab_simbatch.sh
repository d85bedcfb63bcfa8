#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-sb}
timeout 600 python -m pytest tests/test_gpu_similar.py -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; echo "similar tests rc=$?"; tail -5 $OUT/${TAG}_tests.log
for nq in 64 128; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches_q$nq.csv python tools/probe_simbatch.py $nq > $OUT/${TAG}_launches_q$nq.log 2>&1; echo "launches nq=$nq rc=$?"
done
CHECK_BATCH=1 NO_CPU=1 timeout 600 python tools/bench_similar.py > $OUT/${TAG}_bench.log 2>&1; echo "bench rc=$?"; grep "queries\|identical" $OUT/${TAG}_bench.log
