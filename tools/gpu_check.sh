#!/bin/bash
# One GPU-box pass: (optional clean rebuild) parity tests, the default bench line, optional ncu captures.
# Outputs under gpurun_out/<tag>_*.   usage: tools/gpu_check.sh <tag> [tests|bench|ncu|rebuild ...]
tag=${1:-chk}; shift
what=${*:-tests bench}
mkdir -p gpurun_out
for w in $what; do
case $w in
rebuild)  # a clean-checkout nvcc build on the GPU box itself
  ( time python -c "import __graft_entry__ as e; e.build(force=True)" ) > gpurun_out/${tag}_rebuild.log 2>&1; echo "rebuild exit $?"; tail -3 gpurun_out/${tag}_rebuild.log;;
tests)
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/${tag}_tests.log
  tail -5 gpurun_out/${tag}_tests.log;;
bench)
  timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
  python tools/show_bench.py gpurun_out/${tag}_bench.json;;
ncu)
  for wl in cfg2 cfg3 cfg5; do
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${tag}_launches_$wl.csv \
      python bench.py --workload $wl --steps 1 --warmup 3 --no-cpu --no-sub > gpurun_out/${tag}_ncu_list_$wl.log 2>&1
    echo "ncu list $wl exit $?"
  done;;
esac
done
du -sh gpurun_out
