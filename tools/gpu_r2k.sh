#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py tests/test_gpu_kernels.py -m gpu -q -x > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2k_pytest.log
tail -12 gpurun_out/r2k_pytest.log
for args in "109 cfg2 6=1" "112 cfg2 6=1" "106 cfg2 6=1" "109 cfg2 " "106 cfg2 "; do
timeout 120 python tools/ffn_trace.py $args 2>&1 | grep -v Warn | tee -a gpurun_out/r2k_trace.log
done
timeout 300 python tools/ab_knobs.py cfg2 "6=1" "" "7=4" "7=3" 2>&1 | grep -v Warning | tee -a gpurun_out/r2k_ab.log
timeout 300 python tools/ab_knobs.py cfg1 "6=1" "" "7=1" "7=4" 2>&1 | grep -v Warning | tee -a gpurun_out/r2k_ab.log
timeout 300 python tools/ab_knobs.py cfg4_b8 "6=1" "" "7=4" 2>&1 | grep -v Warning | tee -a gpurun_out/r2k_ab.log
