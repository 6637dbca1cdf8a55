#!/bin/bash
# First-contact diagnostics on the B200 box: each group runs in its own process (a CUDA fault poisons the
# context) under its own timeout; everything is logged to gpurun_out/.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
run() { # name, timeout, cmd...
  local name=$1 to=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout "$to" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "rc=$rc $(tail -n 1 gpurun_out/$name.log)" | tee -a gpurun_out/summary.txt
  return $rc
}
PT="python -m pytest -q -m gpu -p no:cacheprovider --timeout 300"
run ints 600 $PT tests/test_gpu_kernels.py -k "integers or durations or prosody"
run attn_f32 600 $PT tests/test_gpu_kernels.py -k "attention and fp32"
run parity_f32 1200 $PT tests/test_gpu_parity.py -k "fp32 or external or composition"
run tc_gemm 600 $PT tests/test_gpu_kernels.py -k "tc_gemm"
if [ $? -ne 0 ]; then
  DX_DEBUG="0=1" run tc_gemm_swap 600 $PT tests/test_gpu_kernels.py -k "tc_gemm"
fi
run attn_bf16 600 $PT tests/test_gpu_kernels.py -k "attention and bf16"
if [ $? -ne 0 ]; then
  DX_DEBUG="1=1" run attn_bf16_swap 600 $PT tests/test_gpu_kernels.py -k "attention and bf16"
fi
run parity_bf16 1200 $PT tests/test_gpu_parity.py -k "bf16"
run smoke 600 python __graft_entry__.py smoke
cat gpurun_out/summary.txt
