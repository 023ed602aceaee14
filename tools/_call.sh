timeout 200 python tools/step_time.py cfg2 --variants v2 --reps 7 2>&1 | grep -v Warn | tail -1
CAPDEC_NO_FFN2_FUSE=1 timeout 200 python tools/step_time.py cfg2 --variants v2 --reps 7 2>&1 | tail -1 | sed 's/^/nofuse /'
timeout 200 python tools/step_time.py cfg2 --variants v2 --reps 7 --images 128 2>&1 | tail -1 | sed 's/^/128img /'
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
