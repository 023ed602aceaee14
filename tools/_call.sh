timeout 300 python bench.py --impl torch_gpu --workload cfg2 --steps 3 --warmup 1 2>/dev/null | tee gpurun_out/torch_gpu_cfg2.json | cut -c1-250
timeout 600 python bench.py --impl torch_gpu --workload cfg3 --steps 1 --warmup 1 2>/dev/null | tee gpurun_out/torch_gpu_cfg3.json | cut -c1-250
timeout 300 python bench.py --impl torch_gpu --workload vit_greedy --images 8 --steps 3 --warmup 1 2>/dev/null | tee gpurun_out/torch_gpu_vit8.json | cut -c1-250
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tee gpurun_out/ref_cfg2.json | cut -c1-300
