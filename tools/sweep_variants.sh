#!/bin/bash
# developer sweep of CM_SCORE_VARIANT / CM_FLAGS on the GPU box
for f in 1; do for v in 8 40 41 42; do echo "== variant $v flags $f"; CM_FLAGS=$f CM_SCORE_VARIANT=$v python tools/probe_kernel.py 4096x16 512x8 65536x16 16384x64 4096x2 reps=300 2>&1 | tail -5; done; done
