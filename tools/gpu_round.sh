#!/bin/bash
# Developer script for one gpurun call: tests, bench, launch list, full captures.  Writes to gpurun_out/.
# usage: tools/gpu_round.sh TAG [steps...]   steps: tests bench launches ncu1 ncu16 similar
set -u
TAG=$1; shift
OUT=gpurun_out
mkdir -p $OUT
for step in "$@"; do
  case $step in
    tests)
      timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?" ;;
    bench)
      timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?" ;;
    bench20)
      timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/${TAG}_bench20.json 2> $OUT/${TAG}_bench20.err; echo "bench20 rc=$?" ;;
    launches)
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches.csv \
        python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-other-shapes --roofline-s 0.002 --min-wall-s 0.001 > $OUT/${TAG}_launches.log 2>&1; echo "launches rc=$?" ;;
    ncu1)
      for shape in 65536x1 4096x1; do
        timeout 600 ncu --set full --clock-control none --import-source on -k regex:cm_score -s 40 -c 1 -f -o $OUT/${TAG}_single_${shape} \
          python tools/probe_kernel.py $shape reps=60 > $OUT/${TAG}_ncu_${shape}.log 2>&1; echo "ncu $shape rc=$?"
      done ;;
    ncu16)
      CM_FLAGS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:cm_score -s 40 -c 1 -f -o $OUT/${TAG}_4096x16 \
          python tools/probe_kernel.py 4096x16 reps=60 > $OUT/${TAG}_ncu_4096x16.log 2>&1; echo "ncu 4096x16 rc=$?" ;;
    similar)
      CM_SIMILAR_GENERIC_SCAN=1 NO_CPU=1 timeout 300 python tools/bench_similar.py > $OUT/${TAG}_similar_generic.log 2>&1
      NO_CPU=1 timeout 300 python tools/bench_similar.py > $OUT/${TAG}_similar_default.log 2>&1
      tail -n 3 $OUT/${TAG}_similar_*.log ;;
    probe)
      CM_FLAGS=1 timeout 600 python tools/probe_kernel.py 4096x16 65536x1 4096x1 512x8 65536x16 reps=2000 > $OUT/${TAG}_probe.log 2>&1; tail -n 8 $OUT/${TAG}_probe.log ;;
    *) echo "unknown step $step" ;;
  esac
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
echo done
