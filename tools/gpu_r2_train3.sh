#!/bin/bash
mkdir -p gpurun_out
timeout 700 python -m pytest tests/test_gpu_train.py tests/test_gpu_kernels.py -q -x > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2u_pytest.log
DX_NO_REF=1 timeout 300 python tools/train_bench.py 32 bf16 > gpurun_out/r2u_train_bench.txt 2>&1; echo "train_bench rc=$?"; head -22 gpurun_out/r2u_train_bench.txt
