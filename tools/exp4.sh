#!/bin/bash
for m in 2049 1024; do
for wl in cfg5; do CAPDEC_WS_MIN_M=$m timeout 200 python tools/step_time.py $wl --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/WS_MIN_M=$m /"; done
CAPDEC_WS_MIN_M=$m timeout 200 python tools/step_time.py cfg3 --images 256 --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/WS_MIN_M=$m cfg4n2 /"
done
CAPDEC_WS_MIN_M=1024 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "full_size" 2>&1 | tail -3
