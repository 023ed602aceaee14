#!/bin/bash
# N-GPU pass: the 2-rank NCCL parity test and the default (cfg4 strong-sharded) bench line with its sub-records.
n=${1:-2}; tag=${2:-mg}; mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two_gpus" > gpurun_out/${tag}_nccl_test.log 2>&1; echo "nccl test exit $?"; tail -3 gpurun_out/${tag}_nccl_test.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n$n.json 2> gpurun_out/${tag}_bench_n$n.err
echo "bench N=$n exit $?"; tail -5 gpurun_out/${tag}_bench_n$n.err; python tools/show_bench.py gpurun_out/${tag}_bench_n$n.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $n --steps 1 --warmup 1 > gpurun_out/${tag}_ref_n$n.json 2> gpurun_out/${tag}_ref_n$n.err
echo "reference arm N=$n exit $?"; cut -c1-400 gpurun_out/${tag}_ref_n$n.json
