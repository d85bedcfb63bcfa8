#!/bin/bash
# developer sweep of CM_SCORE_VARIANT / CM_FLAGS on the GPU box
for f in 1; do for v in 20 30 31 32; do echo "== variant $v flags $f"; CM_FLAGS=$f CM_SCORE_VARIANT=$v python tools/probe_kernel.py 4096x1 65536x1 1000000x1 reps=300 2>&1 | tail -3; done; done
