#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/one_step.py > gpurun_out/r2y_one_step.log 2>&1 && \
timeout 500 ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:attention_tc_kernel" -c 12 -f -o gpurun_out/r2y_attn python tools/one_step.py > gpurun_out/r2y_ncu_attn.log 2>&1
echo "ncu attn rc=$?"; grep -v '^==PROF== Profiling' gpurun_out/r2y_ncu_attn.log | tail -8
