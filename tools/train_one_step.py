"""One training step (B = 32, cfg5 shape) for profilers: warm-up steps, then ONE step between cudaProfilerStart/Stop.
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv ... python tools/train_one_step.py [bf16|fp32]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200 import synthetic  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200.train import TrainEngine  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "bf16"
hp = default_hparams()
batch = tuple(t.cuda() for t in synthetic.forward_batch(hp, batch=32, lmin=20, lmax=150, seed=42))
eng = TrainEngine(hp, make_state_dict(hp, 1234), gemm=mode)
for it in range(1, 4):
    eng.step(batch, iteration=it)
torch.cuda.synchronize()
torch.cuda.profiler.start()
eng.step(batch, iteration=4)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("mode", mode, "loss", float(eng.losses[0]))
