"""Bring-up: does dx_train_step read workspace memory it did not write?  Fill the workspace with NaN bit patterns (0xFF),
run one forward + backward, report the loss terms and every gradient tensor that is not finite, then compare all
gradients with those of a run on a zeroed workspace (must be bit-identical)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200 import synthetic  # noqa: E402
from daft_exprt_b200.train import TrainEngine  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
hp = default_hparams()
sd = make_state_dict(hp, 1234)
batch = tuple(t.cuda() for t in synthetic.forward_batch(hp, batch=B, lmin=10, lmax=40, seed=29))
res = {}
for fill in (0, 0xFF, 0x7F):
    eng = TrainEngine(hp, sd)
    ws = eng._workspace(batch[0].shape[0], batch[0].shape[1], batch[8].shape[2])
    ws.fill_(fill)
    losses = eng.forward_backward(batch, iteration=1)
    torch.cuda.synchronize()
    g = {k: v.clone() for k, v in eng.grads_state_dict().items()}
    res[fill] = (losses.clone(), g)
    bad = [k for k, v in g.items() if not torch.isfinite(v).all()]
    print(f"fill 0x{fill:02X}: losses {losses.tolist()}")
    print(f"   non-finite gradients: {len(bad)}", bad[:12])
for fill in (0xFF, 0x7F):
    diff = [k for k in res[0][1] if not torch.equal(res[0][1][k], res[fill][1][k])]
    print(f"fill 0x{fill:02X} vs zeros: {len(diff)} gradient tensors differ", diff[:12])
    for k in diff[:6]:
        d = (res[0][1][k] - res[fill][1][k]).abs()
        print("   ", k, "max diff", d.max().item(), "n", int((d > 0).sum()), "of", d.numel())
