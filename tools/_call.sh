for w in cfg2 cfg5 cfg3; do for v in base new base new; do
if [ $v = base ]; then export CAPDEC_LIB=$PWD/build/lib_base.so; else unset CAPDEC_LIB; fi
timeout 200 python tools/step_time.py $w --variants v2 --reps 7 2>&1 | tail -1 | sed "s/^/$v /"
done; done
unset CAPDEC_LIB
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
