python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_size or many_row or teacher_forced or golden" 2>&1 | tail -2
python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1
python tools/step_time.py cfg5 --variants v2 --reps 5 2>&1 | tail -1
