timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 200 python tools/handoff_time.py 2>&1 | tail -2
