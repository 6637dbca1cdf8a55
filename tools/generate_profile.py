"""Where the host time of the synthesis driver goes: cProfile of generate.generate_mel_specs (256 utterances, batches of 64).
    python tools/generate_profile.py"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200 import generate as G, synthetic  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.model import DaftExprt  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402

hp = default_hparams()
model = DaftExprt(hp, mode="bf16").cuda()
model.load_state_dict(make_state_dict(hp, 1234))
n_utt, bs = 256, 64
req = synthetic.synthesis_request(hp, n_utt, 20, 150, 42, "add")
gkw = dict(batch_size=bs, alpha_dur=1.0, alpha_pitch=1.0, alpha_energy=1.0)
for _ in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    G.generate_mel_specs(model, output_dir=None, hparams=hp, **req, **gkw)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
print(f"{n_utt} utterances in {dt * 1e3:.2f} ms: {n_utt / dt:.0f} utterances/s, {dt * 1e3 / (n_utt / bs):.3f} ms per batch")
pr = cProfile.Profile()
pr.enable()
G.generate_mel_specs(model, output_dir=None, hparams=hp, **req, **gkw)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
