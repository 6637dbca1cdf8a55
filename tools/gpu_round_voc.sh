#!/bin/bash
# full GPU suite + smoke + bench + vocoder ncu (launch list of one pass, full capture of the conv kernel)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 6 gpurun_out/smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
python tools/voc_one.py 400 16 1 > gpurun_out/voc_one.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv -k regex:vc_ --log-file gpurun_out/voc_launches.csv python tools/voc_one.py 400 16 1 > gpurun_out/ncu_voc_launches.log 2>&1
echo "ncu voc launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:vc_conv_tc --launch-skip 20 --launch-count 57 -f -o gpurun_out/voc_prof python tools/voc_one.py 400 16 1 > gpurun_out/ncu_voc_full.log 2>&1
echo "ncu voc full rc=$?"; tail -n 2 gpurun_out/ncu_voc_full.log
