#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export DX_DEBUG="2=1"
timeout 300 python tools/voc_diag.py 100 4 > gpurun_out/voc_diag_sync.log 2>&1; echo "diag(sync) rc=$?"; grep "rel-L1\|Error\|error\|timeout" gpurun_out/voc_diag_sync.log | head
unset DX_DEBUG
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider --timeout 300 tests/test_gpu_vocoder.py > gpurun_out/voc_tests.log 2>&1; echo "tests rc=$?"; tail -n 15 gpurun_out/voc_tests.log
timeout 300 python tools/voc_diag.py 400 16 > gpurun_out/voc_diag.log 2>&1; echo "diag rc=$?"; grep -A22 "^bf16: B=16" gpurun_out/voc_diag.log
