#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "tile_widths or full_size" > gpurun_out/r2m_tests.log 2>&1; echo "tests exit $?"; tail -15 gpurun_out/r2m_tests.log | cut -c1-200
for v in 1 0; do CAPDEC_NO_WS=$v timeout 200 python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/NO_WS=$v /"; done
