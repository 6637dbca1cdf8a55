"""One cfg2 forward step for profilers: warm-up steps, then ONE step between cudaProfilerStart/Stop.

    python tools/one_step.py [--mode bf16|fp32] [--config cfg2]
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv ... python tools/one_step.py
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200 import synthetic, _lib  # noqa: E402
from daft_exprt_b200.model import DaftExprt  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="bf16")
ap.add_argument("--config", default="cfg2")
ap.add_argument("--warmup", type=int, default=3)
args = ap.parse_args()
hp = default_hparams()
model = DaftExprt(hp, mode=args.mode).cuda()
model.load_state_dict(make_state_dict(hp, 1234))
batch = tuple(t.cuda() for t in synthetic.config_batch(hp, args.config))
for _ in range(args.warmup):
    model(batch)
torch.cuda.synchronize()
lib = _lib.load()
l0 = lib.dx_launch_count()
torch.cuda.profiler.start()
out = model(batch)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("launches", lib.dx_launch_count() - l0, "frames", int(batch[9].sum()), "mel", tuple(out[3][0].shape))
