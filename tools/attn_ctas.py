"""Per-CTA timeline of one attention launch (globaltimer): start / end / SM / K-V tiles of every working CTA.

    python tools/attn_ctas.py [launch 1..12]     # default 12 = decoder block 3 (frame stream)
"""
import os
import sys
import collections

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
batch = tuple(t.cuda() for t in synthetic.config_batch(hp, "cfg2"))
for _ in range(3):
    model(batch)
torch.cuda.synchronize()
lib = _lib.load()
NCTA = 6 * 2 * 64
buf = torch.zeros(256 + 4 * NCTA, dtype=torch.long, device="cuda")
buf[250] = -(1 << 62)
lib.dx_set_trace_buffer(buf.data_ptr(), buf.numel() * 8)
lib.dx_set_debug(10, int(sys.argv[1]) if len(sys.argv) > 1 else 12)
model(batch)
torch.cuda.synchronize()
lib.dx_set_trace_buffer(None, 0)
t = buf.cpu().tolist()
t0 = -t[250]
rows = []
for i in range(NCTA):
    s, e, sm, nkv = t[256 + 4 * i: 260 + 4 * i]
    if s:
        rows.append((i, s - t0, e - t0, sm, nkv))
print(f"{len(rows)} working CTAs, span {(t[251] - t0) / 1000:.2f} us")
by_nkv = collections.defaultdict(list)
for i, s, e, sm, nkv in rows:
    by_nkv[nkv].append(e - s)
print("K/V tiles : CTAs, mean duration ns, ns per tile")
for nkv in sorted(by_nkv):
    d = by_nkv[nkv]
    print(f"  {nkv:3d} : {len(d):4d}  {sum(d) / len(d):8.0f}  {sum(d) / len(d) / nkv:7.0f}")
per_sm = collections.defaultdict(list)
for i, s, e, sm, nkv in rows:
    per_sm[sm].append((s, e, nkv, i))
print("start-time histogram (us): ", collections.Counter(int(s / 2000) * 2 for _, s, _, _, _ in rows))
print("end-time histogram (us):   ", collections.Counter(int(e / 2000) * 2 for _, _, e, _, _ in rows))
busiest = sorted(per_sm.items(), key=lambda kv: -max(e for _, e, _, _ in kv[1]))[:3]
for sm, l in busiest:
    print(f"SM {sm}: " + "  ".join(f"[cta {i} {s / 1000:.1f}-{e / 1000:.1f} us, {n} tiles]" for s, e, n, i in sorted(l)))
print("CTAs per SM:", collections.Counter(len(v) for v in per_sm.values()))
