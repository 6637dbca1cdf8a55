#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/parity_measured.json
python -m pytest tests -m gpu -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -8 gpurun_out/r2b_pytest.log
