#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/train_bench.py 32 bf16 2>&1 | grep -v Warn | head -40 | tee gpurun_out/r2n_train_bench_bf16.log
timeout 600 python -m pytest tests/test_gpu_train.py -m gpu -q 2>&1 | tail -5 | tee gpurun_out/r2n_pytest_train.log
