python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_size or cross_attention or grouped" 2>&1 | tail -3
for v in 1 0 1 0; do echo "== NO_CROSS_RING=$v"; CAPDEC_NO_CROSS_RING=$v python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1; done
for v in 1 0; do echo "== NO_CROSS_RING=$v 256 images"; CAPDEC_NO_CROSS_RING=$v python tools/step_time.py cfg3 --variants v2 --reps 5 --images 256 2>&1 | tail -1; done
