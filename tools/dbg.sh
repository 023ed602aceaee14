for v in 1 0 1 0; do echo "== cfg3 NO_CA_COMPACT=$v"; CAPDEC_NO_CA_COMPACT=$v python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1; done
for v in 1 0; do echo "== cfg3 256 images NO_CA_COMPACT=$v"; CAPDEC_NO_CA_COMPACT=$v python tools/step_time.py cfg3 --variants v2 --reps 5 --images 256 2>&1 | tail -1; done
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_size or cross_attention or grouped or beam or tile_widths" 2>&1 | tail -2
