python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1
python tools/step_time.py cfg5 --variants v2 --reps 5 2>&1 | tail -1
python tools/step_time.py cfg2 --variants v2 --reps 5 2>&1 | tail -1
python tools/full_time.py cfg3 2>&1 | grep "mode=full"
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "many_row or full_size or split_k or tile_widths" 2>&1 | tail -2
