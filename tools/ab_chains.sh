#!/bin/bash
# Developer A/B: build-on-box with -DCM_SPLIT_CHAINS=0/1, probe the launched kernels and the staged one.
OUT=gpurun_out; mkdir -p $OUT
TAG=${1:-ch}
for var in 0 1; do
  CM_NVCC_EXTRA="-DCM_SPLIT_CHAINS=$var" python -m cassmantle_b200.build --force > $OUT/${TAG}_build$var.log 2>&1 || { echo "build $var failed"; tail -5 $OUT/${TAG}_build$var.log; exit 1; }
  CM_FLAGS=1 timeout 300 python tools/probe_kernel.py 4096x16 65536x16 65536x1 16384x1 1000000x1 reps=2000 > $OUT/${TAG}_v${var}_base.log 2>&1
  for c in ${CFGS:-0 1 5}; do
    CM_SCORE_KERNEL=staged CM_STAGED_CFG=$c CM_FLAGS=1 timeout 300 python tools/probe_kernel.py 65536x1 1000000x1 reps=2000 > $OUT/${TAG}_v${var}_cfg$c.log 2>&1
  done
done
grep -H "B=" $OUT/${TAG}_v*.log
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 $OUT/${TAG}_tests.log
