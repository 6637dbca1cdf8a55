#!/bin/bash
mkdir -p gpurun_out
for args in "109 cfg2 6=1" "112 cfg2 6=1" "9 cfg2 6=1" "12 cfg2 6=1" "106 cfg2 6=1"; do
timeout 120 python tools/ffn_trace.py $args 2>&1 | grep -v Warn | tee -a gpurun_out/r2h_trace.log
done
timeout 300 python tools/ab_knobs.py cfg2 "6=1" "" 2>&1 | grep -v Warning | tee -a gpurun_out/r2h_ab.log
