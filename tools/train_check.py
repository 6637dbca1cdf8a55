"""Bring-up: gradients of dx_train_step against the reference's autograd (live reference class on the CPU), per tensor."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402
from daft_exprt_b200 import synthetic  # noqa: E402
from daft_exprt_b200.train import TrainEngine, loss_dict  # noqa: E402
from oracle import train_oracle as TO  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    lmin, lmax = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (5, 30)
    it = 1234
    hp = default_hparams()
    sd = make_state_dict(hp, 1234)
    batch = synthetic.forward_batch(hp, batch=B, lmin=lmin, lmax=lmax, seed=17)
    t0 = time.time()
    ref_loss, ref_ind, ref_grads = TO.reference_step(hp, sd, batch, it)
    print(f"reference fwd+bwd on CPU: {time.time() - t0:.2f} s; loss {ref_loss:.6f} {ref_ind}")
    eng = TrainEngine(hp, sd)
    losses = eng.forward_backward(tuple(t.cuda() for t in batch), iteration=it)
    torch.cuda.synchronize()
    print("ours:", loss_dict(losses))
    g = eng.grads_state_dict()
    worst = 0.0
    for name, r in ref_grads.items():
        o = g[name].cpu()
        d = (o - r).abs()
        rel = d.sum().item() / max(r.abs().sum().item(), 1e-30)
        flag = "" if (d.max().item() <= 1e-3 and rel <= 5e-3) else "   <<<<"
        worst = max(worst, d.max().item())
        print(f"{name:75s} max|ref| {r.abs().max().item():9.3e} max-abs {d.max().item():9.3e} rel-L1 {rel:9.3e}{flag}")
    print("worst max-abs", worst)


if __name__ == "__main__":
    main()
