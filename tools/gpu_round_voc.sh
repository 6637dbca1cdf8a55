#!/bin/bash
# full GPU suite + smoke + bench + vocoder ncu (launch list of one pass, full captures of the fused-layer and conv kernels)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 6 gpurun_out/smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
python tools/voc_one.py 400 16 1 > gpurun_out/voc_one.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv -k regex:vc_ --log-file gpurun_out/voc_launches.csv python tools/voc_one.py 400 16 1 > gpurun_out/ncu_voc_launches.log 2>&1
echo "ncu voc launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:vc_pair --launch-skip 6 --launch-count 6 -f -o gpurun_out/voc_pair python tools/voc_one.py 400 16 1 > gpurun_out/ncu_voc_pair.log 2>&1
echo "ncu pair rc=$?"; tail -n 2 gpurun_out/ncu_voc_pair.log
ncu --set full --clock-control none --import-source on -k regex:vc_conv_tc --launch-skip 26 --launch-count 4 -f -o gpurun_out/voc_s1 python tools/voc_one.py 400 16 1 > gpurun_out/ncu_voc_s1.log 2>&1
echo "ncu s1 rc=$?"
ls -la gpurun_out | head -30
for a in "32 7" "64 7" "64 11"; do python tools/voc_trace.py $a > gpurun_out/trace_${a// /_}.txt 2>&1; done
head -3 gpurun_out/trace_64_7.txt
