"""A/B: fp32 vs bf16 residual stream (HiFiGANGenerator.set_residual) — parity numbers and time."""
import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from oracle import hifigan_oracle as H
import make_vocoder_golden as VG
from daft_exprt_b200 import vocoder as V, _lib
lib = _lib.load()
sd = H.make_state_dict(seed=VG.WEIGHT_SEED)
gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG); gen.load_state_dict(sd); gen.cuda().set_mode("bf16")
mels_t, ln_t = VG.vocoder_case(lengths=[400] * 16, seed=1); mels_t = mels_t.cuda()
cases = {n: (VG.vocoder_case(**kw), np.load(os.path.join(ROOT, "tests", "golden", n + ".npz"))["audio"]) for n, kw in VG.VOCODER_CASES.items()}
wide = VG.vocoder_case(lengths=[120, 77, 64, 9], seed=33)
sd2 = H.make_state_dict(seed=99); ref_wide = H.infer_batch(sd2, *wide).numpy()
gen2 = V.HiFiGANGenerator(V.DEFAULT_CONFIG); gen2.load_state_dict(sd2); gen2.cuda().set_mode("bf16")
for knob in (0, 1):
    gen.set_residual('bf16' if knob else 'fp32'); gen2.set_residual('bf16' if knob else 'fp32')
    for n, ((mels, ln), g) in cases.items():
        a = gen(mels.cuda(), ln)[:, 0].cpu().numpy()
        print(f"x_bf16={knob} {n}: rel-L1 {np.abs(a - g).sum() / np.abs(g).sum():.4e} max-abs {np.abs(a - g).max():.3e}")
    a = gen2(wide[0].cuda(), wide[1])[:, 0].cpu().numpy()
    print(f"x_bf16={knob} wide: rel-L1 {np.abs(a - ref_wide).sum() / np.abs(ref_wide).sum():.4e}")
    for _ in range(3): gen(mels_t, ln_t)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): gen(mels_t, ln_t)
    e1.record(); torch.cuda.synchronize()
    print(f"x_bf16={knob}: {e0.elapsed_time(e1) / 5:.3f} ms per 16 x 400 frames", flush=True)
