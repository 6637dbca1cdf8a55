#!/bin/bash
mkdir -p gpurun_out
timeout 700 python -m pytest tests/test_gpu_train.py -q > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2d_pytest.log
tail -6 gpurun_out/r2d_pytest.log
timeout 300 python - <<'PY' 2>&1 | tee gpurun_out/r2d_step_ms.txt
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
