"""Bring-up of the tensor-core GEMM mode of the training step (TrainEngine(gemm="bf16")): per-tensor gradient difference
against the fp32 mode on the same batch (both on the GPU), loss terms, poisoned-workspace check, timing of both modes."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200 import synthetic  # noqa: E402
from daft_exprt_b200.train import TrainEngine, loss_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
lmin, lmax = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (10, 60)
hp = default_hparams()
sd = make_state_dict(hp, 1234)
batch = tuple(t.cuda() for t in synthetic.forward_batch(hp, batch=B, lmin=lmin, lmax=lmax, seed=29))
res = {}
for mode in ("fp32", "bf16"):
    eng = TrainEngine(hp, sd, gemm=mode)
    eng._workspace(batch[0].shape[0], batch[0].shape[1], batch[8].shape[2]).fill_(0xFF)
    losses = eng.forward_backward(batch, iteration=1000)
    torch.cuda.synchronize()
    res[mode] = (loss_dict(losses), {k: v.clone() for k, v in eng.grads_state_dict().items()})
    print(mode, res[mode][0])
    for it in range(3):
        eng.step(batch, iteration=it + 1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for it in range(5):
        eng.step(batch, iteration=it + 4)
    e1.record()
    torch.cuda.synchronize()
    print(f"{mode}: step {e0.elapsed_time(e1) / 5:.3f} ms")
worst = []
for k, r in res["fp32"][1].items():
    o = res["bf16"][1][k]
    d = (o - r).abs()
    rel = d.sum().item() / max(r.abs().sum().item(), 1e-30)
    worst.append((rel, d.max().item(), r.abs().max().item(), k, bool(torch.isfinite(o).all())))
worst.sort(reverse=True)
print("worst relative L1 (bf16-GEMM mode vs fp32 mode):")
for rel, mx, ref, k, fin in worst[:25]:
    print(f"  {k:75s} rel-L1 {rel:9.3e} max-abs {mx:9.3e} max|ref| {ref:9.3e} finite {fin}")
print("tensors above 2e-2 rel-L1:", sum(1 for w in worst if w[0] > 2e-2), "of", len(worst), " non-finite:", sum(1 for w in worst if not w[4]))
