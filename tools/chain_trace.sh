#!/bin/bash
tag=${1:-tr}; mkdir -p gpurun_out
for wl in ${WLS:-cfg2}; do for cs in ${CSS:-8}; do
  CAPDEC_TRACE_CHAIN=1 CAPDEC_CHAIN_CS=$cs timeout 200 python tools/step_time.py $wl --variants v2 --reps 2 > gpurun_out/${tag}_${wl}_cs$cs.log 2>&1
  echo "== $wl cs=$cs"; tail -8 gpurun_out/${tag}_${wl}_cs$cs.log | cut -c1-700
done; done
