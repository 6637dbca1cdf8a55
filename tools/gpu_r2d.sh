#!/bin/bash
# round 2, call d: the allocation-free entry point's test alone (fresh process), then the whole GPU suite, then the default bench
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2d_zz_alone.log 2>&1; echo "zz alone rc=$?" | tee -a gpurun_out/r2d_zz_alone.log
tail -5 gpurun_out/r2d_zz_alone.log
python -m pytest tests -m gpu -q --durations=15 > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2d_pytest.log
tail -30 gpurun_out/r2d_pytest.log
python bench.py > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/r2d_bench.err
python -c "
import json
d=json.loads(open('gpurun_out/r2d_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'])
print(json.dumps(d.get('configs',{}).get('cfg5'),indent=1)[:3000])
"
