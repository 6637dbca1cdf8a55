#!/bin/bash
mkdir -p gpurun_out
timeout 400 python tools/attn_stress.py 12 2000 > gpurun_out/r2s_attn_stress.txt 2>&1; echo "stress rc=$?"; tail -4 gpurun_out/r2s_attn_stress.txt
timeout 300 python tools/train_bench.py 32 bf16 > gpurun_out/r2s_train_bench.txt 2>&1; echo "train_bench rc=$?"; head -60 gpurun_out/r2s_train_bench.txt
