#!/bin/bash
# head-per-group attention kernel + fused dropout: full GPU suite, train step times, default bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --durations=5 > gpurun_out/r2y_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2y_pytest.log
tail -15 gpurun_out/r2y_pytest.log
timeout 300 python - <<'PY' 2>&1 | tee gpurun_out/r2y_step_ms.txt
import sys, torch
sys.path.insert(0, '.')
from daft_exprt_b200 import synthetic
from daft_exprt_b200.hparams import default_hparams
from daft_exprt_b200.weights import make_state_dict
from daft_exprt_b200.train import TrainEngine
hp = default_hparams()
batch = tuple(t.cuda() for t in synthetic.forward_batch(hp, batch=32, lmin=20, lmax=150, seed=42))
for tm in (False, True):
    eng = TrainEngine(hp, make_state_dict(hp, 1234), gemm="bf16", train_mode=tm)
    for it in range(1, 6): eng.step(batch, iteration=it)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for it in range(6, 26): eng.step(batch, iteration=it)
    e1.record(); torch.cuda.synchronize()
    print("train_mode", tm, "ms/step", e0.elapsed_time(e1) / 20, "loss", float(eng.losses[0]))
PY
timeout 900 python bench.py > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2y_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'])
for k in d['roofline']['kernels']: print(k['kernel'], k['ms_per_step'], k.get('frac'))
c=d['configs']; print('cfg1', c['cfg1']['forward_ms'], c['cfg1']['p50_ms'], 'cfg4', c['cfg4']['b1_ms'], c['cfg4']['b8_ms'], 'cfg3', c['cfg3']['frames_s'], 'cfg5', c['cfg5']['step_ms'], c['cfg5']['eval_mode']['step_ms'], c['cfg5']['fp32_mode']['step_ms'], c['cfg5'].get('eager_gpu_reference',{}).get('step_ms'))
PY
