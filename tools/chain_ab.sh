#!/bin/bash
# A/B of the projection-chain step against the one-kernel-per-projection step (tuning aid).
tag=${1:-ab}; mkdir -p gpurun_out
export CAPDEC_CHAIN_INFO=1
for cs in ${CSS:-8 4}; do
  CAPDEC_CHAIN_CS=$cs timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "rollout or tile_widths or golden or teacher or beam or head_rows" > gpurun_out/${tag}_tests_cs$cs.log 2>&1
  echo "tests cs=$cs exit $?"; tail -4 gpurun_out/${tag}_tests_cs$cs.log
done
for wl in ${WLS:-cfg2 cfg3 cfg5}; do
  [ -z "$NOBASE" ] && CAPDEC_CHAIN=0 timeout 200 python tools/step_time.py $wl --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/nochain /"
  for cs in ${CSS:-8 4}; do for bnb in ${BNBS:-32 64}; do
    CAPDEC_CHAIN_CS=$cs CAPDEC_CHAIN_BNB=$bnb timeout 200 python tools/step_time.py $wl --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/cs=$cs bnb=$bnb /"
  done; done
done
if [ -n "$TRACE" ]; then
for wl in ${TWLS:-cfg2}; do for cs in ${TCSS:-8}; do
  CAPDEC_TRACE_CHAIN=1 CAPDEC_CHAIN_CS=$cs timeout 200 python tools/step_time.py $wl --variants v2 --reps 2 > gpurun_out/${tag}_trace_${wl}_cs$cs.log 2>&1
  echo "== trace $wl cs=$cs"; tail -6 gpurun_out/${tag}_trace_${wl}_cs$cs.log | cut -c1-800
done; done
fi
