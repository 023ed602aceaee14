#!/bin/bash
# part of one decode step (mid-rollout) of cfg3 and one whole step of cfg2 under ncu --set full; only the CSV summaries
# come back (the reports stay in /tmp on the box: gpurun_out/ is limited to 64 MiB)
mkdir -p gpurun_out
python bench.py --workload cfg3 --steps 1 --warmup 3 --no-cpu --no-sub > gpurun_out/r2x_plain_cfg3.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --launch-skip 9575 --launch-count 20 -f -o /tmp/r02_full_cfg3 \
  python bench.py --workload cfg3 --steps 1 --warmup 3 --no-cpu --no-sub > gpurun_out/r2x_ncu_cfg3.log 2>&1
echo "ncu cfg3 exit $?"
python tools/ncu_summary.py /tmp/r02_full_cfg3.ncu-rep > gpurun_out/r02_ncu_full_cfg3.csv
python bench.py --workload cfg2 --steps 1 --warmup 3 --no-cpu --no-sub > gpurun_out/r2x_plain_cfg2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --launch-skip 4580 --launch-count 27 -f -o /tmp/r02_full_cfg2 \
  python bench.py --workload cfg2 --steps 1 --warmup 3 --no-cpu --no-sub > gpurun_out/r2x_ncu_cfg2.log 2>&1
echo "ncu cfg2 exit $?"
python tools/ncu_summary.py /tmp/r02_full_cfg2.ncu-rep > gpurun_out/r02_ncu_full_cfg2.csv
ls -la gpurun_out | tail -8
