#!/bin/bash
# vocoder ncu evidence: launch list of one pass + full captures of conv launches in stage 1 (128 ch) and stage 3 (32 ch)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python tools/voc_one.py 400 16 1 > gpurun_out/voc_one.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv -k regex:vc_ --log-file gpurun_out/voc_launches.csv python tools/voc_one.py 400 16 1 > gpurun_out/ncu_voc_launches.log 2>&1
echo "ncu voc launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:vc_conv_tc --launch-skip 26 --launch-count 6 -f -o gpurun_out/voc_s1 python tools/voc_one.py 400 16 1 > gpurun_out/ncu_voc_s1.log 2>&1
echo "ncu s1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:vc_conv_tc --launch-skip 64 --launch-count 6 -f -o gpurun_out/voc_s3 python tools/voc_one.py 400 16 1 > gpurun_out/ncu_voc_s3.log 2>&1
echo "ncu s3 rc=$?"
ls -la gpurun_out
