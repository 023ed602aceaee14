#!/bin/bash
# One GPU-box pass: parity tests, bench lines for the three single-GPU workloads, ncu launch list and one
# full-set capture of a whole decode step (cfg2).  Outputs under gpurun_out/<tag>_*.
tag=${1:-chk}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${tag}_tests.log
tail -3 gpurun_out/${tag}_tests.log
for w in cfg2 cfg3 cfg5; do
  timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err
  echo "bench $w exit $?"
  python tools/show_bench.py gpurun_out/${tag}_bench_$w.json
done
if [ -z "$NO_NCU" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/${tag}_launches_cfg2.csv \
  python bench.py --workload cfg2 --steps 1 --warmup 3 --no-cpu > gpurun_out/${tag}_ncu_list.log 2>&1
echo "ncu list exit $?"
for w in cfg2 cfg3; do
timeout 900 ncu --set full --clock-control none --launch-skip 620 -c 56 -o /tmp/${tag}_full_$w -f \
  python bench.py --workload $w --steps 1 --warmup 3 --no-cpu > gpurun_out/${tag}_ncu_full_$w.log 2>&1
echo "ncu full $w exit $?"
python tools/ncu_summary.py /tmp/${tag}_full_$w.ncu-rep > gpurun_out/${tag}_ncu_full_$w.csv
done
du -sh gpurun_out
fi
