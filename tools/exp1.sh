#!/bin/bash
mkdir -p gpurun_out
for b in 128 64; do CAPDEC_BNB=$b timeout 200 python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1 | sed "s/^/BNB=$b /"; done
for wl in cfg3 cfg2; do
python bench.py --workload $wl --steps 1 --warmup 3 --no-cpu --no-sub > gpurun_out/r2l_plain_$wl.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2l_launches_$wl.csv \
  python bench.py --workload $wl --steps 1 --warmup 3 --no-cpu --no-sub > gpurun_out/r2l_ncu_list_$wl.log 2>&1
echo "ncu list $wl exit $?"
done
ls -la gpurun_out | tail -8
