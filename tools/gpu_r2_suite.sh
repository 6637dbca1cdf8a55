#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_arch.py -q > gpurun_out/r2v_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2v_pytest.log
tail -4 gpurun_out/r2v_pytest.log
