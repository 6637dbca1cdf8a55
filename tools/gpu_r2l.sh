#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py tests/test_gpu_kernels.py -m gpu -q > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2l_pytest.log
tail -4 gpurun_out/r2l_pytest.log
for args in "109 cfg2 6=1" "112 cfg2 6=1" "106 cfg2 6=1"; do
timeout 120 python tools/ffn_trace.py $args 2>&1 | grep -v Warn | tee -a gpurun_out/r2l_trace.log
done
timeout 300 python tools/ab_knobs.py cfg2 "6=1" "6=1,15=16" "" "7=4" 2>&1 | grep -v Warning | tee -a gpurun_out/r2l_ab.log
timeout 300 python tools/ab_knobs.py cfg1 "6=1" "" "7=4" 2>&1 | grep -v Warning | tee -a gpurun_out/r2l_ab.log
timeout 300 python tools/ab_knobs.py cfg4_b8 "6=1" "" "7=4" 2>&1 | grep -v Warning | tee -a gpurun_out/r2l_ab.log
