"""CPU: the oracle port (oracle/daft_oracle.py) against the golden vectors frozen from the live
reference (tests/golden/make_golden.py).  Travels to the GPU box (no /root/reference needed)."""
import pytest
import torch

import _cases as C
from oracle import daft_oracle as O

FP_TOL = 2e-5   # fp32 CPU vs fp32 CPU, different op grouping allowed


@pytest.mark.parametrize("name", sorted(C.MG.FORWARD_CASES))
def test_forward_matches_golden(name):
    hp, batch, g = C.forward_case(name)
    sd = C.state_dict()
    spk, film, enc, dec, w = O.forward(sd, hp, tuple(x.clone() for x in batch))
    assert torch.equal(dec[1], C.t(g["output_lengths"]))
    assert (spk - C.t(g["speaker_preds"])).abs().max() <= FP_TOL
    assert (film[3] - C.t(g["film_dec"])).abs().max() <= FP_TOL
    assert (dec[0] - C.t(g["mel"])).abs().max() <= FP_TOL
    assert (w - C.t(g["weights"])).abs().max() <= FP_TOL
    acc = O.accent_encoder(sd, hp, batch[6], batch[7], batch[8], batch[9])
    assert (acc - C.t(g["accent_emb"])).abs().max() <= FP_TOL


@pytest.mark.parametrize("name", sorted(C.MG.INFERENCE_CASES))
def test_inference_matches_golden(name):
    hp, pt, inputs, prosody, ext_emb, acc_emb, g = C.inference_case(name)
    sd = C.state_dict()
    enc, dec, w = O.inference(sd, hp, tuple(x.clone() for x in inputs), pt,
                              {k: v.clone() for k, v in prosody.items()}, ext_emb, acc_emb)
    # integer contract: bit-exact
    assert torch.equal(enc[1], C.t(g["durations_int"]))
    assert torch.equal(dec[1], C.t(g["output_lengths"]))
    assert (enc[0] - C.t(g["duration_preds"])).abs().max() == 0
    assert (enc[2] - C.t(g["energy_preds"])).abs().max() <= FP_TOL
    assert (enc[3] - C.t(g["pitch_preds"])).abs().max() <= FP_TOL
    assert (dec[0] - C.t(g["mel"])).abs().max() <= FP_TOL
    assert (w - C.t(g["weights"])).abs().max() <= FP_TOL


def test_length_regulator_integers_consistency():
    hp, batch, g = C.forward_case("fwd_pad")
    ints = O.length_regulator_integers(batch[2], batch[5])
    assert torch.equal(ints["output_lengths"], C.t(g["output_lengths"]))
    assert ints["t_max"] == g["mel"].shape[2]
    # frame→symbol map is monotone and covers exactly d_l frames per symbol
    for b in range(batch[2].shape[0]):
        n = int(ints["output_lengths"][b])
        f2s = ints["frame2symbol"][b, :n]
        assert torch.all(f2s[1:] >= f2s[:-1])
        counts = torch.bincount(f2s, minlength=batch[2].shape[1])
        assert torch.equal(counts, batch[2][b])
        assert torch.all(ints["frame2symbol"][b, n:] == -1)


def test_errors_match_reference_contract():
    hp, batch, _ = C.forward_case("fwd_tiny")
    sd = C.state_dict()
    with pytest.raises(ValueError):
        O.forward(sd, hp, batch[:11])
    inputs, prosody, ext_emb, acc_emb = C.synthetic.inference_batch(hp, 2, 5, 6, 1)
    with pytest.raises(ValueError):
        O.inference(sd, hp, inputs, "add", prosody, None, acc_emb)
    with pytest.raises(ValueError):
        O.inference(sd, hp, inputs, "add", prosody, ext_emb, None)
    with pytest.raises(ValueError):
        O.inference(sd, hp, inputs, "add", None, ext_emb, acc_emb)
    with pytest.raises(NotImplementedError):
        O.inference(sd, hp, inputs, "bogus", prosody, ext_emb, acc_emb)


def test_parse_batch_mirror_equals_reference_golden_on_cpu():
    """DaftExprt.parse_batch (model.py:858-887) is host logic + casts: the mirror on 'cpu' against the frozen output of the
    reference's own method (tests/golden/make_parse_batch_golden.py)."""
    import torch
    import make_parse_batch_golden as MP
    from daft_exprt_b200.model import DaftExprt
    hp, batch14 = MP.parse_batch_case()
    g = C.load_golden("parse_batch")
    inputs, targets = DaftExprt(hp).parse_batch("cpu", batch14)
    for n, t in zip(MP.NAMES_IN, inputs):
        ref = torch.from_numpy(g["in_" + n])
        assert t.dtype == ref.dtype and torch.equal(t, ref), n
    for n, t in zip(MP.NAMES_TGT, targets):
        ref = torch.from_numpy(g["tgt_" + n])
        assert t.dtype == ref.dtype and torch.equal(t, ref), n
