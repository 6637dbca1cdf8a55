#!/bin/bash
mkdir -p gpurun_out
timeout 700 python -m pytest tests/test_gpu_train.py -q -x > gpurun_out/r2t_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2t_pytest.log
timeout 300 python tools/train_bench.py 32 bf16 > gpurun_out/r2t_train_bench.txt 2>&1; echo "train_bench rc=$?"; head -24 gpurun_out/r2t_train_bench.txt
timeout 200 python tools/train_one_step.py bf16 > gpurun_out/r2t_one.log 2>&1 && \
timeout 400 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2t_train_launches.csv python tools/train_one_step.py bf16 > gpurun_out/r2t_ncu.log 2>&1
echo "ncu rc=$?"
