"""Quick A/B of the forward's kernels: cfg2 step time and the per-kernel profile, cfg1 / cfg4 forward times.
    python tools/attn_time.py"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200 import _lib, synthetic  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.model import DaftExprt  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402

hp = default_hparams()
model = DaftExprt(hp, mode="bf16").cuda()
model.load_state_dict(make_state_dict(hp, 1234))
lib = _lib.load()
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")


def timed(batch, n=30):
    for _ in range(5):
        model(batch)
    ms = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        model(batch)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    return ms[len(ms) // 2]


for name in ("cfg2", "cfg1", "cfg4_b1", "cfg4_b8"):
    if name not in synthetic.CONFIGS:
        continue
    batch = tuple(t.cuda() for t in synthetic.config_batch(hp, name))
    print(name, "B", batch[0].shape[0], "forward median ms", round(timed(batch), 4))
    if name == "cfg2":
        lib.dx_profile_enable(1)
        model(batch)
        buf = ctypes.create_string_buffer(1 << 16)
        lib.dx_profile_report(buf, len(buf))
        lib.dx_profile_enable(0)
        print(buf.value.decode())
