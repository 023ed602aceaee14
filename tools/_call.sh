for w in cfg2 cfg5 cfg3; do for v in prev new prev new; do
if [ $v = prev ]; then export CAPDEC_LIB=$PWD/build/lib_prev.so; else unset CAPDEC_LIB; fi
timeout 200 python tools/step_time.py $w --variants v2 --reps 7 2>&1 | tail -1 | sed "s/^/$v /"
done; done
unset CAPDEC_LIB
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
