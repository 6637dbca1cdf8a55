#!/bin/bash
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo8.txt 2>&1; lscpu | grep -E "^CPU\(s\)|NUMA|Model name|Socket" >> gpurun_out/r2_topo8.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench_8gpu.json 2> gpurun_out/r2_bench_8gpu.err; echo "rc=$?"
tail -c 600 gpurun_out/r2_bench_8gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 --no-bind --no-configs > gpurun_out/r2_bench_8gpu_nobind.json 2> gpurun_out/r2_bench_8gpu_nobind.err; echo "rc=$?"
