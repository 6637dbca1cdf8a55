#!/bin/bash
# round-2 evidence run: full GPU suite, default bench, reference arm (short), ncu launch list of one cfg2 step
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/r2z_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2z_pytest.log
tail -14 gpurun_out/r2z_pytest.log
timeout 900 python bench.py > gpurun_out/r2z_bench.json 2> gpurun_out/r2z_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2z_bench_reference.json 2> gpurun_out/r2z_bench_reference.err; echo "reference arm rc=$?"
timeout 200 python tools/one_step.py > gpurun_out/r2z_one_step.log 2>&1 && \
timeout 400 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2z_launches.csv python tools/one_step.py > gpurun_out/r2z_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2z_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'])
c=d['configs']; print('cfg1', c['cfg1']['forward_ms'], c['cfg1']['p50_ms'], 'cfg4', c['cfg4']['b1_ms'], c['cfg4']['b8_ms'], 'cfg3', c['cfg3']['frames_s'], 'cfg5', c['cfg5']['step_ms'], c['cfg5']['eval_mode']['step_ms'], c['cfg5']['fp32_mode']['step_ms'], c['cfg5'].get('eager_gpu_reference',{}).get('step_ms'))
print('vocoder', d['vocoder']['ms_per_batch'], 'cpu_baseline', d['cpu_baseline'])
PY
