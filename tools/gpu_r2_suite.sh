#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/r2v_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2v_pytest.log
tail -25 gpurun_out/r2v_pytest.log
