"""CPU: the mask-injection hook of the training oracle (oracle/train_oracle.py) on the unmodified reference model.

The engine's train-mode step is checked on the GPU against the reference run with the SAME dropout masks; that only means
something if the masks reach the reference's dropout calls in the order and with the shapes the engine assumes
(train.TrainEngine.dropout_masks).  This test pins that order on the reference itself: 3 pre-net sites, then for each stack
(accent encoder, phoneme encoder, frame decoder) and block: attention weights, attention output, feed-forward output."""
import pytest
import torch

import _cases as C
from oracle import ref_harness as R
from oracle import train_oracle as TO

pytestmark = pytest.mark.skipif(not R.reference_available(), reason="reference (oracle/_ref) not present")


def _masks(B, L, T, p, gen, keep_all=False):
    def m(shape):
        return (torch.ones(shape, dtype=torch.bool) if keep_all else torch.rand(shape, generator=gen) >= p, p)
    out = [m((B, T, 1024)), m((B, T, 1024)), m((B, T, 128))]
    for n in (T, L, T):
        for _ in range(4):
            out += [m((B * 2, n, n)), m((B, n, 128)), m((B, n, 128))]
    return out


def test_reference_reaches_its_dropout_sites_in_the_order_the_engine_exports_masks():
    hp = C.hparams()
    sd = C.state_dict()
    batch = C.synthetic.forward_batch(hp, batch=3, lmin=5, lmax=20, seed=5)
    B, L = batch[0].shape
    T = batch[8].shape[2]
    gen = torch.Generator().manual_seed(1)
    masks = _masks(B, L, T, 0.1, gen)
    loss_eval, _, _ = TO.reference_step(hp, sd, batch, 100)
    loss_a, _, g_a = TO.reference_step(hp, sd, batch, 100, dropout_masks=masks)      # asserts shapes, p and the call count
    loss_b, _, g_b = TO.reference_step(hp, sd, batch, 100, dropout_masks=masks)
    assert loss_a == loss_b and all(torch.equal(g_a[k], g_b[k]) for k in g_a), "injected masks make train mode reproducible"
    assert loss_a != loss_eval
    with pytest.raises(AssertionError):
        TO.reference_step(hp, sd, batch, 100, dropout_masks=masks[:-1])                # one mask short
    with pytest.raises(AssertionError):
        TO.reference_step(hp, sd, batch, 100, dropout_masks=masks[1:] + masks[:1])     # wrong order: shapes disagree
    import torch.nn.functional as F
    assert F.dropout.__module__ == "torch.nn.functional", "the stock dropout is restored afterwards"
