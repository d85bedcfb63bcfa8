#!/bin/bash
# Developer A/B: the staged one-word kernel (CM_SCORE_KERNEL=staged, CM_STAGED_CFG=k) against the shipped ones.
OUT=gpurun_out; mkdir -p $OUT
TAG=${1:-st}
CM_SCORE_KERNEL=staged timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $OUT/${TAG}_parity_cfg0.log 2>&1; echo "parity cfg0 rc=$?"
CM_SCORE_KERNEL=staged CM_STAGED_CFG=2 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $OUT/${TAG}_parity_cfg2.log 2>&1; echo "parity cfg2 rc=$?"
SHAPES="65536x1 16384x1 262144x1 1000000x1"
CM_FLAGS=1 timeout 300 python tools/probe_kernel.py $SHAPES reps=2000 > $OUT/${TAG}_base.log 2>&1
for c in ${CFGS:-0 1 2 3 4 5 6 7}; do
  CM_SCORE_KERNEL=staged CM_STAGED_CFG=$c CM_FLAGS=1 timeout 300 python tools/probe_kernel.py $SHAPES reps=2000 > $OUT/${TAG}_cfg$c.log 2>&1
done
grep -H "B=" $OUT/${TAG}_base.log $OUT/${TAG}_cfg*.log
