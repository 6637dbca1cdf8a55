"""A/B timing of debug-knob settings on the forward pass: python tools/ab_knobs.py cfg2 "" "6=1" "7=1" "7=3,8=3" ...
Prints ms per step (CUDA events, L2 flushed between steps, median of N) and the per-kernel profile of the block tails."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200 import synthetic, _lib  # noqa: E402
from daft_exprt_b200.model import DaftExprt  # noqa: E402

cfg = sys.argv[1]
settings = sys.argv[2:] or [""]
hp = default_hparams()
model = DaftExprt(hp, mode="bf16").cuda()
model.load_state_dict(make_state_dict(hp, 1234))
batch = tuple(t.cuda() for t in synthetic.config_batch(hp, cfg))
lib = _lib.load()
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
import ctypes


def run(setting, n=30):
    knobs = [tuple(int(x) for x in kv.split("=")) for kv in setting.split(",") if kv]
    for k, v in knobs:
        lib.dx_set_debug(k, v)
    try:
        for _ in range(5):
            model(batch)
        torch.cuda.synchronize()
        ts = []
        for _ in range(n):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            model(batch)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        lib.dx_profile_enable(1)
        for _ in range(5):
            model(batch)
        buf = ctypes.create_string_buffer(1 << 16)
        lib.dx_profile_report(buf, len(buf))
        lib.dx_profile_enable(0)
        prof = {}
        for line in buf.value.decode().splitlines():
            tag, cnt, ms = line.split()
            prof[tag] = float(ms) / 5
        tails = {k: round(v, 4) for k, v in prof.items() if "block_tail" in k or k in ("attn_tc",)}
        print(f"{cfg} knobs[{setting or 'default'}]: median {ts[len(ts) // 2]:.4f} ms  min {ts[0]:.4f}  {tails}", flush=True)
    finally:
        for k, _ in knobs:
            lib.dx_set_debug(k, 0)


for s in settings:
    run(s)
