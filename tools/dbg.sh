python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tile_widths or full_size or many_row or split_k or full_sequence" 2>&1 | tail -3
for v in 1 0 1 0; do echo "== cfg3 NO_PAIR_K=$v"; CAPDEC_NO_PAIR_K=$v python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1; done
for v in 1 0; do echo "== full cfg3 NO_PAIR_K=$v"; CAPDEC_NO_PAIR_K=$v python tools/full_time.py cfg3 2>&1 | grep "mode=full"; done
