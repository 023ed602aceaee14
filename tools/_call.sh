bash tools/gpu_check.sh r1h
