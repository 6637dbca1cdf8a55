#!/bin/bash
# A/B: GPU parity tests, then bench with the default path and with the debug knobs given as arguments ("7=1" ...)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest -q -m gpu -x -p no:cacheprovider --timeout 300 tests > gpurun_out/ab_tests.log 2>&1; echo "tests rc=$?"; tail -n 5 gpurun_out/ab_tests.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ab_default.json 2> gpurun_out/ab_default.err; echo "default rc=$?"; cat gpurun_out/ab_default.json
for k in "$@"; do
  DX_DEBUG="$k" timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > "gpurun_out/ab_$k.json" 2> "gpurun_out/ab_$k.err"; echo "knob $k rc=$?"; cat "gpurun_out/ab_$k.json"
done
