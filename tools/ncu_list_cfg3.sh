#!/bin/bash
# ncu launch list (gpu__time_duration only) of the default bench command's workload, after the plain run exited 0
mkdir -p gpurun_out
python bench.py --workload cfg3 --steps 1 --warmup 3 --no-cpu --no-sub --no-verify > gpurun_out/r2z_plain.log 2>&1 || { echo plain failed; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file /tmp/r02_launches_cfg3.csv \
  python bench.py --workload cfg3 --steps 1 --warmup 3 --no-cpu --no-sub --no-verify > gpurun_out/r2z_ncu.log 2>&1
echo "ncu exit $?"
python tools/launch_summary.py /tmp/r02_launches_cfg3.csv > gpurun_out/r02_launch_summary_cfg3.csv
wc -l /tmp/r02_launches_cfg3.csv gpurun_out/r02_launch_summary_cfg3.csv
