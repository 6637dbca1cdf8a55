"""Does L2 blocking pay?  The same 16 x 400-frame batch vocoded as 1, 2, 4, 8, 16 generator passes."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_vocoder_golden as VG
from daft_exprt_b200 import vocoder as V
gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG); gen.load_state_dict(V.synthetic_state_dict()); gen.cuda().set_mode("bf16")
B, T = 16, 400
mels, ln = VG.vocoder_case(lengths=[T] * B, seed=1); mels = mels.cuda()
full = gen(mels, ln)
for groups in (1, 2, 4, 8, 16):
    g = B // groups
    parts = [(mels[i * g:(i + 1) * g].contiguous(), ln[i * g:(i + 1) * g]) for i in range(groups)]
    def run():
        return [gen(m, l) for m, l in parts]
    for _ in range(2): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): out = run()
    e1.record(); torch.cuda.synchronize()
    same = torch.equal(torch.cat(out), full)
    print(f"{groups:2d} passes of {g * T} frames: {e0.elapsed_time(e1) / 5:.3f} ms, identical {same}", flush=True)
