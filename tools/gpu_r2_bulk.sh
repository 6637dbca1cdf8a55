#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -k "bulk or fine_tune or packed" > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2x_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2x_bench.json').read().strip().splitlines()[-1])
c=d['configs']; print('value', d['value'], 'e2e', d['e2e']['value'], 'cfg3', json.dumps(c['cfg3'])[:600])
PY
