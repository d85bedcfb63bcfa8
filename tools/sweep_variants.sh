#!/bin/bash
# developer sweep of CM_SCORE_VARIANT / CM_FLAGS on the GPU box
for f in 1; do for v in 8; do echo "== variant $v flags $f"; CM_FLAGS=$f CM_SCORE_VARIANT=$v python tools/probe_kernel.py 4096x16 512x8 65536x16 16384x64 4096x2 65536x2 reps=300 2>&1 | tail -6; done; done
for f in 0 1; do for v in 20 21 22 23 24 25; do echo "== variant $v flags $f"; CM_FLAGS=$f CM_SCORE_VARIANT=$v python tools/probe_kernel.py 4096x1 65536x1 reps=300 2>&1 | tail -2; done; done
