for wl in cfg3 cfg5; do for v in 8 4 2 8 4; do echo "== $wl SELF_ATTN_ITS=$v"; CAPDEC_SELF_ATTN_ITS=$v python tools/step_time.py $wl --variants v2 --reps 5 2>&1 | tail -1; done; done
CAPDEC_SELF_ATTN_ITS=4 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_size or greedy_rollout or grouped" 2>&1 | tail -2
