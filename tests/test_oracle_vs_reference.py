"""CPU, build container only: the oracle port against the LIVE reference imported from
/root/reference (skipped where the reference is absent, e.g. on the GPU box)."""
import copy
import pytest
import torch

import _cases as C
from oracle import daft_oracle as O
from oracle import ref_harness as R

pytestmark = pytest.mark.skipif(not R.reference_available(), reason="/root/reference not present")


@pytest.fixture(scope="module")
def ref():
    return R.build_reference_model(C.hparams(), C.state_dict())


def test_state_dict_layout_is_the_references(ref):
    sd = C.state_dict()
    rsd = ref.state_dict()
    assert list(rsd.keys()) == list(sd.keys())
    assert len(sd) == 184
    assert sum(v.numel() for v in sd.values()) == 14385891
    for k in sd:
        assert tuple(rsd[k].shape) == tuple(sd[k].shape), k


@pytest.mark.parametrize("kw", [dict(batch=2, lmin=3, lmax=9, seed=21), dict(batch=5, lmin=10, lmax=30, seed=22),
                                dict(batch=2, lmin=30, lmax=30, seed=23, const_dur=4)])
def test_forward_port_equals_reference(ref, kw):
    hp = C.hparams()
    batch = C.synthetic.forward_batch(hp, **kw)
    with torch.no_grad():
        r = ref(tuple(x.clone() for x in batch))
    o = O.forward(C.state_dict(), hp, tuple(x.clone() for x in batch))
    assert torch.equal(r[3][1], o[3][1])
    for a, b in ((r[0], o[0]), (r[1][3], o[1][3]), (r[3][0], o[3][0]), (r[4], o[4])):
        assert (a - b).abs().max() <= 2e-5
    # integer side of the length regulator (what the reference computes inline, model.py:485-497)
    ints = O.length_regulator_integers(batch[2], batch[5])
    assert torch.equal(ints["cumsum"], torch.cumsum(batch[2], dim=1))
    assert ints["t_max"] == r[4].shape[2]


@pytest.mark.parametrize("pt", ["add", "multiply"])
def test_inference_port_equals_reference(ref, pt):
    hp = C.hparams()
    inputs, prosody, ee, ae = C.synthetic.inference_batch(hp, 4, 6, 25, 31, pt)
    with torch.no_grad():
        r = ref.inference(tuple(x.clone() for x in inputs), pt, hp, {k: v.clone() for k, v in prosody.items()},
                          ee.clone(), ae.clone())
    o = O.inference(C.state_dict(), hp, tuple(x.clone() for x in inputs), pt,
                    {k: v.clone() for k, v in prosody.items()}, ee.clone(), ae.clone())
    assert torch.equal(r[0][1], o[0][1])          # durations_int
    assert torch.equal(r[1][1], o[1][1])          # output_lengths
    assert torch.equal(r[0][0], o[0][0])          # thresholded float durations
    for a, b in ((r[0][2], o[0][2]), (r[0][3], o[0][3]), (r[1][0], o[1][0]), (r[2], o[2])):
        assert (a - b).abs().max() <= 2e-5


def test_int_durations_edge_cases(ref):
    """Short symbols dropped (< filter_length/sr/2), zero rows, and the reference's error paths."""
    hp = C.hparams()
    g = torch.Generator().manual_seed(5)
    for trial in range(20):
        B, L = 3, 17
        hops = torch.rand(B, L, generator=g) * 6
        dur = hops * (hp.hop_length / hp.sampling_rate)
        dur[0, 3] = 0.001          # below threshold → zeroed
        dur[1, -4:] = 0.0          # trailing padding
        try:
            rd, ri = ref.get_int_durations(dur.clone(), hp)
            r_err = None
        except Exception as e:     # reference can fail on degenerate rows
            r_err = type(e)
        try:
            od, oi = O.get_int_durations(dur.clone(), hp)
            o_err = None
        except Exception as e:
            o_err = type(e)
        assert (r_err is None) == (o_err is None), (trial, r_err, o_err)
        if r_err is None:
            assert torch.equal(ri, oi)
            assert torch.equal(rd, od)
