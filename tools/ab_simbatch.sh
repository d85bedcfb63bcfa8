#!/bin/bash
# Developer run: batched most_similar -- tests, timing, launch lists and full captures of the tensor-core scan.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-sb}
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 $OUT/${TAG}_tests.log
CM_SIMILAR_PAIR=1 timeout 300 python -m pytest tests/test_gpu_similar.py -m gpu -x -q -k "batch or many" > $OUT/${TAG}_tests_pair.log 2>&1; echo "similar tests (pair forced) rc=$?"; tail -2 $OUT/${TAG}_tests_pair.log
CM_SIMILAR_PAIR=0 timeout 300 python -m pytest tests/test_gpu_similar.py -m gpu -x -q -k "batch or many" > $OUT/${TAG}_tests_single.log 2>&1; echo "similar tests (single forced) rc=$?"; tail -2 $OUT/${TAG}_tests_single.log
CHECK_BATCH=1 NO_CPU=1 timeout 300 python tools/bench_similar.py > $OUT/${TAG}_bench.log 2>&1; echo "bench rc=$?"; grep "queries\|identical\|p50" $OUT/${TAG}_bench.log
for nq in 64 256; do
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches_q$nq.csv python tools/probe_simbatch.py $nq > $OUT/${TAG}_launches_q$nq.log 2>&1; echo "launches nq=$nq rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cm_simscan -s 3 -c 1 -f -o $OUT/${TAG}_simscan_q$nq python tools/probe_simbatch.py $nq > $OUT/${TAG}_ncu_q$nq.log 2>&1; echo "ncu nq=$nq rc=$?"
done
