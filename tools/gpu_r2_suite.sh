#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/arch_grad_debug.py 2>&1 | grep -v "Warning\|re.findall" > gpurun_out/r2v_debug.txt; tail -5 gpurun_out/r2v_debug.txt
