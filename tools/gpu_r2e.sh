#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py tests/test_gpu_kernels.py -m gpu -q -x > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2e_pytest.log
tail -25 gpurun_out/r2e_pytest.log
for cfg in cfg2 cfg1 cfg4_b8; do
timeout 300 python tools/ab_knobs.py $cfg "" "6=1" "7=1" "7=3" "7=4" "8=1" "8=4" "7=1,8=1" 2>&1 | grep -v Warning | tee -a gpurun_out/r2e_ab.log
done
