#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_train.py -q -x > gpurun_out/r2c_pytest_train.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest_train.log
tail -15 gpurun_out/r2c_pytest_train.log
python tools/train_bench.py 32 > gpurun_out/r2c_train_bench.log 2>&1; grep -v Warning gpurun_out/r2c_train_bench.log | tail -45
