for w in cfg2 cfg5; do for p in 1 2 4 1 2 4; do
CAPDEC_WHOLE_X_PARTS=$p timeout 200 python tools/step_time.py $w --variants v2 --reps 7 2>&1 | tail -1 | sed "s/^/parts=$p /"
done; done
CAPDEC_WHOLE_X_PARTS=2 timeout 600 python -m pytest tests -m gpu -x -q -k "tile_widths or golden or greedy_rollout or grouped_rollout" 2>&1 | tail -2
CAPDEC_WHOLE_X_PARTS=4 timeout 600 python -m pytest tests -m gpu -x -q -k "tile_widths or golden or greedy_rollout or grouped_rollout" 2>&1 | tail -2
