python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "graph_and_direct or full_size or beam or batch_invariant or greedy_rollout or teacher_forced" 2>&1 | tail -2
for v in 4 0 8 0; do echo "== cfg3 SELF_ATTN_ITS=$v (0 = by step)"; CAPDEC_SELF_ATTN_ITS=$v python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1; done
for v in 4 0; do echo "== cfg5 SELF_ATTN_ITS=$v"; CAPDEC_SELF_ATTN_ITS=$v python tools/step_time.py cfg5 --variants v2 --reps 5 2>&1 | tail -1; done
