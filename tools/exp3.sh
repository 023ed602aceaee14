#!/bin/bash
for d in 0 1 2 3; do
CAPDEC_WS_DBG=$d python bench.py --workload cfg3 --steps 2 --warmup 3 --no-cpu --no-sub > gpurun_out/r2o_$d.json 2>gpurun_out/r2o_$d.err
echo "dbg=$d"; python tools/show_bench.py gpurun_out/r2o_$d.json | grep "gemm_ffn1\|gemm_qkv\|gemm_cross_q\|value="
done
