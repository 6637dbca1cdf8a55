#!/bin/bash
# ncu source-level capture of one frame-stream block tail with P3 (launch 9 of a step), whole tiles per CTA
mkdir -p gpurun_out
export DX_DEBUG="6=1"
timeout 200 python tools/one_step.py > gpurun_out/r2i_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ffn_tc_kernel -s 44 -c 1 -o gpurun_out/r2i_ffn python tools/one_step.py > gpurun_out/r2i_ncu.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/r2i_ncu.log
