"""clock64 timeline of CTA 0 of the fused ResBlock-layer kernel (vc_pair_tc_kernel), via dx_set_trace_buffer.

    python tools/voc_trace.py [C=32|64] [taps=3|7|11]     # the last matching launch of one pass is kept
Columns (cycles since the CTA entered its main loop): producer: A-tile copies issued; issuer: M1 start (D1 free),
M1 first stage landed, M1 issued; M2: Hs ready, D2 free, M2 issued; epilogue warp 0: P1 entered, D1 ready, P1 done,
P2 entered, D2 ready, P2 done."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_vocoder_golden as VG
from daft_exprt_b200 import vocoder as V, _lib
lib = _lib.load()
Cf = int(sys.argv[1]) if len(sys.argv) > 1 else 32
taps = int(sys.argv[2]) if len(sys.argv) > 2 else 7
gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG); gen.load_state_dict(V.synthetic_state_dict()); gen.cuda().set_mode("bf16")
mels, ln = VG.vocoder_case(lengths=[400] * 16, seed=1); mels = mels.cuda()
for _ in range(2): gen(mels, ln)
torch.cuda.synchronize()
buf = torch.zeros(8 + 16 * 64, dtype=torch.long, device="cuda")
lib.dx_set_trace_buffer(buf.data_ptr(), buf.numel() * 8)
lib.dx_set_debug(10, Cf); lib.dx_set_debug(15, taps)
gen(mels, ln); torch.cuda.synchronize()
lib.dx_set_trace_buffer(None, 0); lib.dx_set_debug(10, 0); lib.dx_set_debug(15, 0)
t = buf.cpu().tolist()
t0, t1, n = t[0], t[1], t[2]
print(f"C={Cf} taps={taps}: CTA 0 ran {n} tiles in {t1 - t0} cycles ({(t1 - t0) / max(n, 1):.0f} per tile)")
names = ["copy_iss", "-", "M1_go", "M1_land", "M1_done", "Hs_rdy", "D2_free", "M2_done", "P1_in", "D1_rdy", "P1_done", "P2_in", "D2_rdy", "P2_done"]
print("tile " + " ".join(f"{x:>8s}" for x in names))
for j in range(min(n, 24)):
    row = t[8 + 16 * j: 8 + 16 * j + 14]
    print(f"{j:4d} " + " ".join(f"{(v - t0) if v else 0:8d}" for v in row))
