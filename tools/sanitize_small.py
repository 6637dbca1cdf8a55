"""Small forward (both modes) + one training step (both modes) for compute-sanitizer --tool memcheck."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200 import synthetic  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200.model import DaftExprt  # noqa: E402
from daft_exprt_b200.train import TrainEngine  # noqa: E402

hp = default_hparams()
sd = make_state_dict(hp, 1234)
batch = tuple(t.cuda() for t in synthetic.forward_batch(hp, batch=3, lmin=2, lmax=40, seed=5))
for mode in ("bf16", "fp32"):
    m = DaftExprt(hp, mode=mode).cuda()
    m.load_state_dict(sd)
    out = m(batch)
    torch.cuda.synchronize()
    print("forward", mode, float(out[3][0].abs().sum()))
for gemm in ("bf16", "fp32"):
    eng = TrainEngine(hp, sd, gemm=gemm)
    losses = eng.step(batch, iteration=10)
    torch.cuda.synchronize()
    print("train", gemm, float(losses[0]))
