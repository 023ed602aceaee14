#!/bin/bash
# cfg4 per-GPU shapes: 128 images (N=4, 768 rows) and 64 images (N=8, 384 rows)
for img in 128 64; do
CAPDEC_CHAIN=0 timeout 200 python tools/step_time.py cfg3 --images $img --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/img=$img nochain /"
CAPDEC_CHAIN=0 CAPDEC_WS_MIN_M=256 timeout 200 python tools/step_time.py cfg3 --images $img --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/img=$img nochain ws256 /"
CAPDEC_CHAIN=2 CAPDEC_CHAIN_BNB=64 timeout 200 python tools/step_time.py cfg3 --images $img --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/img=$img chain64 /"
CAPDEC_CHAIN=2 CAPDEC_CHAIN_BNB=32 timeout 200 python tools/step_time.py cfg3 --images $img --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/img=$img chain32 /"
done
