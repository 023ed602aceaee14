#!/bin/bash
# ncu --set full of the LAST decode step of a warm cfg3 rollout (final round-2 kernels); only the CSV summary comes back
mkdir -p gpurun_out
python tools/one_rollout.py cfg3 2 > gpurun_out/r2w_plain.log 2>&1 || { echo plain failed; tail -3 gpurun_out/r2w_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 100000 --csv --log-file /tmp/r2w_list.csv python tools/one_rollout.py cfg3 2 > gpurun_out/r2w_list.log 2>&1
K=$(grep -c '^"' /tmp/r2w_list.csv); K=$((K-1)); echo "kernels in the run: $K"
SKIP=$((K-51)); echo "skip $SKIP"
timeout 1200 ncu --set full --clock-control none --launch-skip $SKIP --launch-count 51 -f -o /tmp/r02_full_cfg3_final python tools/one_rollout.py cfg3 2 > gpurun_out/r2w_ncu.log 2>&1
echo "ncu exit $?"
python tools/ncu_summary.py /tmp/r02_full_cfg3_final.ncu-rep > gpurun_out/r02_ncu_full_cfg3_final.csv
wc -l gpurun_out/r02_ncu_full_cfg3_final.csv
