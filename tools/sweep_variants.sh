#!/bin/bash
# developer sweep of CM_SCORE_VARIANT / CM_FLAGS on the GPU box
for f in 1; do for v in 8 5 6; do echo "== variant $v flags $f"; CM_FLAGS=$f CM_SCORE_VARIANT=$v python tools/probe_kernel.py 4096x16 8192x16 16384x16 65536x16 16384x64 512x8 16384x4 16384x8 reps=300 2>&1 | tail -8; done; done
