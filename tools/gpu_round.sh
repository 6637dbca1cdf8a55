#!/bin/bash
# tests + bench + ncu launch list (+ optional full capture of one kernel: $1 = kernel regex, $2 = launches to capture, default 12)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q -p no:cacheprovider --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
python bench.py --steps 5 --warmup 3 --mode fp32 --no-cpu-baseline > gpurun_out/bench_fp32.json 2>> gpurun_out/bench.err; echo "bench fp32 rc=$?"; cat gpurun_out/bench_fp32.json
python tools/one_step.py > gpurun_out/one_step.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/one_step.py > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"; tail -n 2 gpurun_out/one_step.log
if [ -n "$1" ]; then
  ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:$1" -c ${2:-12} -f -o gpurun_out/prof python tools/one_step.py > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"; tail -n 3 gpurun_out/ncu_full.log
fi
