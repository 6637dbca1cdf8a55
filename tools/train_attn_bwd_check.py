"""Bring-up of the tcgen05 attention backward: tensor-core training mode with the fp32 attention backward (debug knob 16)
against the same mode with dx_attn_bwd_tc.cu, per gradient tensor; then the timing of both."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200 import _lib, synthetic  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200.train import TrainEngine, loss_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
lmin, lmax = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (10, 60)
hp = default_hparams()
sd = make_state_dict(hp, 1234)
batch = tuple(t.cuda() for t in synthetic.forward_batch(hp, batch=B, lmin=lmin, lmax=lmax, seed=29))
lib = _lib.load()
res = {}
for knob in (1, 0):
    lib.dx_set_debug(16, knob)
    eng = TrainEngine(hp, sd, gemm="bf16")
    eng._workspace(batch[0].shape[0], batch[0].shape[1], batch[8].shape[2]).fill_(0xFF)
    losses = eng.forward_backward(batch, iteration=1000)
    torch.cuda.synchronize()
    res[knob] = {k: v.clone() for k, v in eng.grads_state_dict().items()}
    print("fp32 attention backward" if knob else "tcgen05 attention backward", loss_dict(losses)["loss"])
    for it in range(3):
        eng.step(batch, iteration=it + 1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for it in range(5):
        eng.step(batch, iteration=it + 4)
    e1.record()
    torch.cuda.synchronize()
    print(f"   step {e0.elapsed_time(e1) / 5:.3f} ms")
lib.dx_set_debug(16, 0)
worst = []
for k, r in res[1].items():
    o = res[0][k]
    d = (o - r).abs()
    worst.append((d.sum().item() / max(r.abs().sum().item(), 1e-30), d.max().item(), r.abs().max().item(), k, bool(torch.isfinite(o).all())))
worst.sort(reverse=True)
for rel, mx, ref, k, fin in worst[:14]:
    print(f"  {k:75s} rel-L1 {rel:9.3e} max-abs {mx:9.3e} max|ref| {ref:9.3e} finite {fin}")
print("non-finite:", sum(1 for w in worst if not w[4]), " above 5e-2:", sum(1 for w in worst if w[0] > 5e-2))
