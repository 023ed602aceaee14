python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r1c_bench_cfg2_n8.json 2> gpurun_out/r1c_bench_cfg2_n8.err; echo "n8 cfg2 exit $?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --steps 5 --warmup 3 --workload cfg5 > gpurun_out/r1c_bench_cfg5_n8.json 2> gpurun_out/r1c_bench_cfg5_n8.err; echo "n8 cfg5 exit $?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 3 --workload cfg3 > gpurun_out/r1c_bench_cfg3_n8.json 2> gpurun_out/r1c_bench_cfg3_n8.err; echo "n8 cfg3 exit $?"
python tools/show_bench.py gpurun_out/r1c_bench_cfg2_n8.json gpurun_out/r1c_bench_cfg5_n8.json gpurun_out/r1c_bench_cfg3_n8.json | grep value
tail -3 gpurun_out/r1c_bench_cfg2_n8.err
