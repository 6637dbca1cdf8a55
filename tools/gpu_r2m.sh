#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2m_pytest.log
tail -4 gpurun_out/r2m_pytest.log
timeout 300 python tools/ab_knobs.py cfg2 "" "6=1" 2>&1 | grep -v Warning | tee -a gpurun_out/r2m_ab.log
timeout 300 python tools/ab_knobs.py cfg1 "" "6=1" 2>&1 | grep -v Warning | tee -a gpurun_out/r2m_ab.log
timeout 300 python tools/ab_knobs.py cfg4_b8 "" "6=1" 2>&1 | grep -v Warning | tee -a gpurun_out/r2m_ab.log
timeout 600 python bench.py > gpurun_out/r2m_bench.json 2> gpurun_out/r2m_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2m_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'], d['kernels_ms_per_step'])
print(d['configs']['cfg1']['forward_ms'], d['configs']['cfg1']['p50_ms'], d['configs']['cfg4']['b1_ms'], d['configs']['cfg4']['b8_ms'])
"
