"""Timing experiments on the vocoder conv kernel (debug knob 11; results are garbage, only the time matters)."""
import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from oracle import hifigan_oracle as H
import make_vocoder_golden as VG
from daft_exprt_b200 import vocoder as V, _lib
lib = _lib.load()
gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG); gen.load_state_dict(H.make_state_dict(seed=VG.WEIGHT_SEED)); gen.cuda().set_mode("bf16")
mels, ln = VG.vocoder_case(lengths=[400] * 16, seed=1); mels = mels.cuda()
def prof(tagfilter=("vc_s2", "vc_s3", "vc_s1", "vc_s0")):
    lib.dx_profile_enable(1); gen(mels, ln); torch.cuda.synchronize()
    buf = ctypes.create_string_buffer(1 << 16); lib.dx_profile_report(buf, len(buf)); lib.dx_profile_enable(0)
    d = {}
    for l in buf.value.decode().strip().splitlines():
        t, n, ms = l.split(); d[t] = float(ms)
    return d
for k in (0, 1, 2, 8, 16, 17, 24, 25, 27):
    lib.dx_set_debug(11, k)
    gen(mels, ln); torch.cuda.synchronize()
    d = prof()
    print(f"knob {k:2d}: total {sum(d.values()):7.3f} ms | " + " ".join(f"{t[3:]}={d[t]:.2f}" for t in sorted(d) if t.startswith("vc_s")), flush=True)
lib.dx_set_debug(11, 0)
