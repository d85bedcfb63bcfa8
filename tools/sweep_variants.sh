#!/bin/bash
# developer sweep of CM_SCORE_VARIANT / CM_FLAGS on the GPU box
for f in 0 1; do for v in 1 8 10 11 12 13 14; do echo "== variant $v flags $f"; CM_FLAGS=$f CM_SCORE_VARIANT=$v python tools/probe_kernel.py 4096x16 512x8 65536x16 16384x64 reps=300 2>&1 | tail -4; done; done
for f in 0 1; do for v in 3 4 5 6 7; do echo "== variant $v flags $f"; CM_FLAGS=$f CM_SCORE_VARIANT=$v python tools/probe_kernel.py 4096x1 65536x1 reps=300 2>&1 | tail -2; done; done
