#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
PT="python -m pytest -q -m gpu -p no:cacheprovider --timeout 600"
export DX_DEBUG="2=1"
timeout 1500 $PT tests/test_gpu_parity.py -k "fp32 or external or composition" > gpurun_out/parity_f32.log 2>&1; echo "f32 rc=$?"; tail -n 3 gpurun_out/parity_f32.log
timeout 1500 $PT tests/test_gpu_parity.py -k "bf16" > gpurun_out/parity_bf16.log 2>&1; echo "bf16 rc=$?"; tail -n 3 gpurun_out/parity_bf16.log
unset DX_DEBUG
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 4 gpurun_out/smoke.log
