#!/bin/bash
# round-2 additions under ncu --set full: the prefill on the weight-stationary GEMM and the full-sequence pass (first
# two layers), cfg3 shapes with 128 images (one chunk of 37 632 rows); only the CSV summary comes back
mkdir -p gpurun_out
python tools/one_full_pass.py cfg3 --images 128 > gpurun_out/r2y_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r2y_plain.log; exit 1; }
timeout 1200 ncu --set full --clock-control none --profile-from-start off --launch-count 22 -f -o /tmp/r02_full_fullpass \
  python tools/one_full_pass.py cfg3 --images 128 > gpurun_out/r2y_ncu.log 2>&1
echo "ncu exit $?"
python tools/ncu_summary.py /tmp/r02_full_fullpass.ncu-rep > gpurun_out/r02_ncu_full_fullpass.csv
wc -l gpurun_out/r02_ncu_full_fullpass.csv
