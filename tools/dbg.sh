for v in 96 97 96 97; do echo "== WS_MIN_CTAS=$v"; CAPDEC_WS_MIN_CTAS=$v python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1; done
