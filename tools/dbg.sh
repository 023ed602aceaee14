run() { n=$1; shift; echo "== images=$n $*"; env "$@" python tools/step_time.py cfg3 --variants v2 --reps 5 --images $n 2>&1 | tail -1; }
run 128 A=1
run 128 CAPDEC_BNB=64
run 128 CAPDEC_CHAIN=2
run 128 CAPDEC_WS_MIN_CTAS=64
run 256 A=1
run 256 CAPDEC_BNB=128
run 256 CAPDEC_BNB=32
