#!/bin/bash
mkdir -p gpurun_out
for args in "109 cfg2 6=1" "109 cfg2 6=1,15=2" "109 cfg2 6=1,15=3" "109 cfg2 6=1,15=15" "109 cfg4_b8 7=4" "109 cfg4_b8 7=4,15=3"  "109 cfg2 " "109 cfg1 " "106 cfg1 "; do
timeout 120 python tools/ffn_trace.py $args 2>&1 | grep -v Warn | tee -a gpurun_out/r2g_trace.log
done
timeout 300 python tools/ab_knobs.py cfg2 "6=1" "6=1,15=1" "6=1,15=2" "6=1,15=3" "6=1,15=4" "6=1,15=8" "6=1,15=15" 2>&1 | grep -v Warning | tee -a gpurun_out/r2g_ab.log
