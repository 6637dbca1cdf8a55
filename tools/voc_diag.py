"""GPU bring-up of the HiFi-GAN generator: parity numbers per mode on the golden cases, then timing + per-kernel
profile of a wider batch.   python tools/voc_diag.py [frames_per_utt] [batch]"""
import ctypes
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from oracle import hifigan_oracle as H
import make_vocoder_golden as VG
from daft_exprt_b200 import vocoder as V, _lib

lib = _lib.load()
sd = H.make_state_dict(seed=VG.WEIGHT_SEED)
for mode in ("fp32", "bf16"):
    gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG)
    gen.load_state_dict(sd)
    gen.cuda().set_mode(mode)
    for name, kw in VG.VOCODER_CASES.items():
        g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))["audio"]
        mels, ln = VG.vocoder_case(**kw)
        a = gen(mels.cuda(), ln)[:, 0].cpu().numpy()
        print(f"{mode} {name}: max-abs {np.abs(a - g).max():.3e} rel-L1 {np.abs(a - g).sum() / np.abs(g).sum():.3e} "
              f"finite {np.isfinite(a).all()}", flush=True)

T = int(sys.argv[1]) if len(sys.argv) > 1 else 400
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
mels, ln = VG.vocoder_case(lengths=[T] * B, seed=1)
mels = mels.cuda()
flops_per_frame = 0
for name, kind, cin, cout, k, sd_ in H.conv_names():
    rate = {512: 1, 256: 8, 128: 64, 64: 128, 32: 256}
    if kind == "conv":
        r = 1 if name == "conv_pre" else rate[cin]
        flops_per_frame += 2 * cin * cout * k * r
    else:
        flops_per_frame += 2 * cin * cout * 2 * rate[cout]
print(f"algorithmic FLOP / frame: {flops_per_frame / 1e6:.1f} M")
for mode in ("bf16", "fp32"):
    if mode == "fp32" and B * T > 2000:
        continue
    gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG)
    gen.load_state_dict(sd)
    gen.cuda().set_mode(mode)
    for _ in range(3):
        gen(mels, ln)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    e0.record()
    for _ in range(n):
        gen(mels, ln)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    fr = B * T
    print(f"{mode}: B={B} T={T}: {ms:.3f} ms/call, {fr / ms * 1e3 / 1e6:.3f} M frames/s, {fr * 256 / ms * 1e3 / 1e6:.1f} M samples/s, "
          f"{fr * flops_per_frame / ms / 1e9:.1f} TFLOP/s", flush=True)
    lib.dx_profile_enable(1)
    gen(mels, ln)
    torch.cuda.synchronize()
    buf = ctypes.create_string_buffer(1 << 16)
    lib.dx_profile_report(buf, len(buf))
    lib.dx_profile_enable(0)
    print(buf.value.decode())
