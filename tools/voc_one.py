"""One bf16 HiFi-GAN forward (for ncu captures).  python tools/voc_one.py [frames_per_utt] [batch] [calls]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from oracle import hifigan_oracle as H
import make_vocoder_golden as VG
from daft_exprt_b200 import vocoder as V

T = int(sys.argv[1]) if len(sys.argv) > 1 else 400
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
calls = int(sys.argv[3]) if len(sys.argv) > 3 else 1
gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG)
gen.load_state_dict(H.make_state_dict(seed=VG.WEIGHT_SEED))
gen.cuda().set_mode("bf16")
mels, ln = VG.vocoder_case(lengths=[T] * B, seed=1)
mels = mels.cuda()
for _ in range(calls):
    a = gen(mels, ln)
torch.cuda.synchronize()
print("ok", tuple(a.shape), float(a.abs().mean()))
