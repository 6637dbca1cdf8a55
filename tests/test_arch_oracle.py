"""CPU: the oracle port on architectures other than the default 4 / 4 / 4 blocks, with an ``hparams.accent_encoder``
sub-dict of its own (model.py:627-631), against golden vectors frozen from the live reference built from the same hparams
(tests/golden/make_arch_golden.py)."""
import pytest
import torch

import _cases as C
from oracle import daft_oracle as O

FP_TOL = 2e-5


@pytest.mark.parametrize("name", sorted(C.MA.ARCH_CASES))
def test_oracle_forward_matches_golden_on_other_block_counts(name):
    hp, sd, batch, inf, g = C.arch_case(name)
    n_ae, n_pe, n_fd = C.MA.ARCH_CASES[name][:3]
    assert sum(k.startswith("accent_encoder.blocks.") and k.endswith("feed_forward.layer_norm.bias") for k in sd) == n_ae
    assert sd["style_adapter.gammas_predictor.linear_layer.weight"].shape[0] == (n_pe + n_fd) * 128
    spk, film, enc, dec, w = O.forward(sd, hp, tuple(x.clone() for x in batch))
    assert film[3].shape == (batch[0].shape[0], n_fd, 256)
    assert torch.equal(dec[1], C.t(g["output_lengths"]))
    assert (spk - C.t(g["speaker_preds"])).abs().max() <= FP_TOL
    assert (film[3] - C.t(g["film_dec"])).abs().max() <= FP_TOL
    assert (dec[0] - C.t(g["mel"])).abs().max() <= FP_TOL
    assert (w - C.t(g["weights"])).abs().max() <= FP_TOL
    acc = O.accent_encoder(sd, hp, batch[6], batch[7], batch[8], batch[9])
    assert (acc - C.t(g["accent_emb"])).abs().max() <= FP_TOL
    if inf is not None:
        pt, inputs, prosody, ext_emb, acc_emb = inf
        enc, dec, w = O.inference(sd, hp, tuple(x.clone() for x in inputs), pt, {k: v.clone() for k, v in prosody.items()},
                                  ext_emb, acc_emb)
        assert torch.equal(enc[1], C.t(g["inf_durations_int"])) and torch.equal(dec[1], C.t(g["inf_output_lengths"]))
        assert (dec[0] - C.t(g["inf_mel"])).abs().max() <= FP_TOL
        assert (w - C.t(g["inf_weights"])).abs().max() <= FP_TOL
