echo "== decode 12544 rows"; python tools/step_time.py cfg2 --variants v2 --reps 1 --images 12544 2>&1 | tail -2
for st in 2 3 5 6; do echo "== full cfg2 STAGES=$st"; CAPDEC_STAGES=$st CAPDEC_FULL_SYNC=1 python tools/full_time.py cfg2 2>&1 | grep -i "capdec error\|mode=full\|max |"; done
