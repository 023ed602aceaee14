for v in 0 16 32 48 0 24; do echo "== cfg3 CA_PREFETCH_MB=$v"; CAPDEC_CA_PREFETCH_MB=$v python tools/step_time.py cfg3 --variants v2 --reps 5 2>&1 | tail -1; done
