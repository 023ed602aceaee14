#!/bin/bash
for img in 128 256 512; do for c in 96 64 32; do
CAPDEC_WS_MIN_CTAS=$c timeout 200 python tools/step_time.py cfg3 --images $img --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/img=$img minctas=$c /"
done; done
for c in 96 64 32; do CAPDEC_WS_MIN_CTAS=$c timeout 200 python tools/step_time.py cfg5 --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/cfg5 minctas=$c /"; done
