#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/generate_profile.py > gpurun_out/r2v_generate_profile.txt 2>&1; head -40 gpurun_out/r2v_generate_profile.txt
timeout 200 python -m pytest tests -m gpu -q -k "bulk or fine_tune or packed or generate" 2>&1 | tail -2
