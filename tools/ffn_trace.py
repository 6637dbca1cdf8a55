"""Timeline of CTA 0 of the fused FFN kernel (clock64 deltas), via dx_set_trace_buffer.

    python tools/ffn_trace.py [launch 1..12] [config] [knobs "6=1,7=2"]    # cycles relative to kernel start, per hidden chunk
"""
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
cfg = sys.argv[2] if len(sys.argv) > 2 else "cfg2"
batch = tuple(t.cuda() for t in synthetic.config_batch(hp, cfg))
for kv in (sys.argv[3].split(",") if len(sys.argv) > 3 and sys.argv[3] else []):
    _lib.load().dx_set_debug(int(kv.split("=")[0]), int(kv.split("=")[1]))
print(f"== {cfg} launch {sys.argv[1] if len(sys.argv) > 1 else 'last'} knobs {sys.argv[3] if len(sys.argv) > 3 else ''}")
for _ in range(3):
    model(batch)
torch.cuda.synchronize()
lib = _lib.load()
buf = torch.zeros(256, dtype=torch.long, device="cuda")
buf[250] = -(1 << 62)
lib.dx_set_trace_buffer(buf.data_ptr(), buf.numel() * 8)
if len(sys.argv) > 1 and int(sys.argv[1]) % 100 > 0:
    lib.dx_set_debug(10, int(sys.argv[1]))      # which of the 12 block-tail launches (1-4 accent, 5-8 symbols, 9-12 decoder)
model(batch)          # every FFN launch overwrites the trace: the last one (decoder block 3, frame stream) remains
torch.cuda.synchronize()
lib.dx_set_trace_buffer(None, 0)
t = buf.cpu().tolist()
t0 = t[0]
if t[5]:
    print(f"P0 A landed     {t[5] - t0:8d}")
    print(f"P0 acc loaded   {t[6] - t0:8d}")
print(f"X ready         {t[1] - t0:8d}")
names = ["d1_empty", "G1_issued", "hs_full0", "G2_issued", "w_full0", "-", "d1_full", "tmem_ld1", "hs_empty1", "Hs_done1"]
print("chunk " + " ".join(f"{n:>9s}" for n in names))
for c in range(8):
    row = [t[8 + 12 * c + i] for i in range(10)]
    print(f"{c:5d} " + " ".join(f"{(v - t0) if v else 0:9d}" for v in row))
if t[120]:
    print(f"reduce start    {t[120] - t0:8d}")
    print(f"reduce done     {t[121] - t0:8d}")
if t[124]:
    print(f"P0 row_add done {t[124] - t0:8d}   stats done {t[125] - t0:8d}")
print(f"E2 start        {t[2] - t0:8d}")
if t[122]:
    print(f"E2 row_add done {t[122] - t0:8d}   stats done {t[123] - t0:8d}")
if t[126]:
    print(f"P3 acc loaded   {t[126] - t0:8d} {t[127] - t0:8d} {t[128] - t0:8d}")
print(f"E2 done         {t[3] - t0:8d}")
if t[7]:
    print(f"P3 done         {t[7] - t0:8d}")
print(f"CTA done        {t[4] - t0:8d}")

if len(sys.argv) > 2:
    sys.exit(0)
if t[251]:
    print(f"attention kernel (last launch): first CTA start to last CTA end {(t[251] + t[250]) / 1000:.2f} us, {t[252]} working CTAs")
print("attention CTA (0,0,0) of the last launch (decoder block 3, longest utterance, head 0, q tile 0)")
a0 = t[128]
print(f"Q landed {t[129] - a0}   end {t[130] - a0}")
an = ["k_full", "S_issued", "p_full", "PV_issued", "s_full", "passA", "P_stored", "K_loaded", "V_loaded"]
print(" step " + " ".join(f"{n:>9s}" for n in an))
for j in range(8):
    row = [t[132 + 10 * j + i] for i in range(9)]
    if not any(row):
        continue
    print(f"{j:5d} " + " ".join(f"{(v - a0) if v else 0:9d}" for v in row))
