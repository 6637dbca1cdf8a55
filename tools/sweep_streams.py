"""A/B sweep on one B200: utterance-group streams x attention fusion, cfg2 bf16, resident inputs.
Prints ms/step and checks that every configuration is bit-identical to the single-stream result."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200 import synthetic, _lib  # noqa: E402
from daft_exprt_b200.model import DaftExprt  # noqa: E402

hp = default_hparams()
model = DaftExprt(hp, mode="bf16").cuda()
model.load_state_dict(make_state_dict(hp, 1234))
batch = tuple(t.cuda() for t in synthetic.config_batch(hp, sys.argv[1] if len(sys.argv) > 1 else "cfg2"))
lib = _lib.load()
flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device="cuda")
ref = None
for knob6 in (0, 1):
    lib.dx_set_debug(6, knob6)
    for streams in (1, 2, 3, 4, 6):
        model.streams = streams
        for _ in range(3):
            out = model(batch)
        torch.cuda.synchronize()
        ts = []
        for _ in range(20):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            out = model(batch)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        mel = out[3][0]
        if ref is None:
            ref = mel.clone()
        same = torch.equal(mel, ref) if knob6 == 0 else float((mel - ref).abs().max())
        ts.sort()
        print(f"fused_attn={knob6} streams={streams}: median {ts[10]:.3f} ms  min {ts[0]:.3f} ms  vs-ref {same}", flush=True)
