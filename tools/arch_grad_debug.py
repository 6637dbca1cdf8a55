import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import _cases as C
from daft_exprt_b200 import synthetic
from daft_exprt_b200.weights import make_state_dict
from daft_exprt_b200.train import TrainEngine, loss_dict
from oracle import train_oracle as TO

def run(tag, hp, sd, batch):
    ref_loss, _, ref = TO.reference_step(hp, sd, batch, 3000)
    eng = TrainEngine(hp, sd)
    l = loss_dict(eng.forward_backward(tuple(t.cuda() for t in batch), iteration=3000))
    g = eng.grads_state_dict()
    bad = []
    for n, r in ref.items():
        d = (g[n].cpu() - r).abs()
        rel = d.sum().item() / max(r.abs().sum().item(), 1e-12)
        if d.max().item() > 1e-3 or rel > 2e-3:
            idx = d.flatten().argmax().item()
            bad.append((n, d.max().item(), rel, r.abs().max().item(), tuple(r.shape), idx))
    print(tag, "loss", l["loss"], ref_loss, "bad params", len(bad), "of", len(ref))
    for b in bad[:12]:
        print("   ", b)

hp, sd, batch, _, _ = C.arch_case("arch_143")
run("arch_143", hp, sd, batch)
hp0 = C.hparams()
run("default arch, same batch kwargs", hp0, C.state_dict(), synthetic.forward_batch(hp0, batch=3, lmin=5, lmax=25, seed=23))
import make_arch_golden as MA
for n_ae in (1, 2, 3):
    hp = C.hparams()
    hp.accent_encoder = dict(hp.phoneme_encoder, nb_blocks=n_ae)
    run(f"ae={n_ae} pe=4 fd=4", hp, make_state_dict(hp, 1234), synthetic.forward_batch(hp, batch=3, lmin=5, lmax=25, seed=23))
hp = C.hparams()
hp.frame_decoder = dict(hp.frame_decoder, nb_blocks=3)
run("ae=4 pe=4 fd=3", hp, make_state_dict(hp, 1234), synthetic.forward_batch(hp, batch=3, lmin=5, lmax=25, seed=23))


def run_tc(tag, hp, sd, batch):
    dev = tuple(t.cuda() for t in batch)
    B, L = batch[0].shape
    out = []
    for tm in (False, True):
        eng = TrainEngine(hp, sd, gemm="bf16", train_mode=tm, seed=5)
        masks = [(m.cpu(), pm) for m, pm in eng.dropout_masks(B, L, batch[8].shape[2], iteration=3000)] if tm else None
        l = loss_dict(eng.forward_backward(dev, iteration=3000))["loss"]
        ref_loss, _, ref = TO.reference_step(hp, sd, batch, 3000, dropout_masks=masks)
        g = eng.grads_state_dict()
        rels = sorted((((g[n].cpu() - r).abs().sum() / r.abs().sum().clamp_min(1e-12)).item(), n) for n, r in ref.items() if r.abs().sum() > 1e-6)
        out.append((tm, abs(l - ref_loss) / abs(ref_loss), rels[len(rels) // 2][0], rels[-1]))
    print(tag, "TC mode (train_mode, loss rel, median rel-L1, worst):", out)


print("---- tensor-core mode")
hp, sd, batch, _, _ = C.arch_case("arch_143")
run_tc("arch_143", hp, sd, batch)
hp, sd, batch, _, _ = C.arch_case("arch_232")
run_tc("arch_232", hp, sd, batch)
hp0 = C.hparams()
run_tc("default arch, arch_143's batch", hp0, C.state_dict(), synthetic.forward_batch(hp0, batch=3, lmin=5, lmax=25, seed=23))
