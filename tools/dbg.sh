python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_size or batch_invariant or beam or full_sequence or tf32" 2>&1 | tail -3
for v in 1 0 1 0; do echo "== NO_TF32_HEAD=$v cfg3"; CAPDEC_NO_TF32_HEAD=$v python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1; done
for v in 1 0; do echo "== NO_TF32_HEAD=$v cfg5"; CAPDEC_NO_TF32_HEAD=$v python tools/step_time.py cfg5 --variants v2 --reps 5 2>&1 | tail -1; done
python tools/full_time.py cfg3 2>&1 | tail -3
