#!/bin/bash
mkdir -p gpurun_out
for args in "12 cfg2 6=1" "9 cfg2 6=1" "12 cfg2 " "6 cfg2 6=1" "6 cfg2 " "12 cfg1 6=1" "12 cfg1 " "9 cfg1 " "6 cfg1 "; do
timeout 120 python tools/ffn_trace.py $args 2>&1 | grep -v Warn | tee -a gpurun_out/r2f_trace.log
done
