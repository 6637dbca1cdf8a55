"""Synthesis driver (reference generate.py:104-399).

CPU: the oracle restatement (oracle/generate_oracle.py) against the frozen reference outputs (tests/golden/gen_*.npz)
and, in the build container, against the live reference; the product's host side (collate, external-prosody tensors,
errors) bit-exact against the oracle / the reference.  GPU: the product's ``generate_mel_specs`` through the CUDA path
against the oracle and the golden files."""
import copy
import os
import sys
import tempfile

import numpy as np
import pytest
import torch

import _cases as C
from daft_exprt_b200 import generate as G
from daft_exprt_b200 import synthetic
from oracle import generate_oracle as GO
from oracle import ref_harness as R

sys.path.insert(0, C.GOLDEN_DIR)
import make_generate_golden as MGG          # noqa: E402

needs_reference = pytest.mark.skipif(not R.reference_available(), reason="/root/reference not present")
SRC = MGG.SOURCE_STATS


def _collate_args(req, pt):
    return (req["sentences"], req["dur_factors"], req["energy_factors"], req["pitch_factors"][1], pt, req["speaker_ids"],
            list(req["file_names"]), C.hparams())


def _same_collate(x, y):
    for k in range(6):
        assert x[k].dtype == y[k].dtype and torch.equal(x[k], y[k]), k
    assert x[6] == y[6]
    assert (x[7] is None) == (y[7] is None)
    if x[7] is not None:
        assert all(p is q for p, q in zip(x[7], y[7]))


# ---------------------------------------------------------------- CPU: oracle pinned ----------------------------------
@pytest.mark.parametrize("name", list(MGG.GENERATE_CASES))
def test_oracle_generate_equals_golden(name):
    rk, gk = MGG.GENERATE_CASES[name]
    hp = C.hparams()
    req = synthetic.synthesis_request(hp, **rk)
    g = C.load_golden(name)
    assert str(g["request_digest"]) == MGG.request_digest(req), "synthetic generator drifted"
    preds = GO.generate(C.state_dict(), hp, **copy.deepcopy(req), **gk)
    assert list(preds) == [str(n) for n in g["names"]]
    for k, vals in preds.items():
        for f, v in zip(MGG.FIELDS, vals):
            ref = g[f"{k}/{f}"]
            assert v.shape == ref.shape, (k, f)
            if f in ("duration", "duration_int"):
                assert np.array_equal(v, ref), (k, f)
            else:
                assert np.abs(v - ref).max() <= 2e-5, (k, f)


@needs_reference
@pytest.mark.parametrize("pt,src,alphas", [("add", None, (1.0, 1.0, 1.0)), ("multiply", SRC, (1.3, 0.7, 1.2))])
def test_oracle_generate_equals_live_reference(pt, src, alphas):
    hp = C.hparams()
    ref = R.build_reference_model(hp, C.state_dict())
    req = synthetic.synthesis_request(hp, 5, 4, 12, 17, pt)
    kw = dict(batch_size=2, source_stats=src, alpha_dur=alphas[0], alpha_pitch=alphas[1], alpha_energy=alphas[2])
    with tempfile.TemporaryDirectory() as tmp:
        r = R.run_reference_generate(ref, hp, tmp, **copy.deepcopy(req), **kw)
        assert sorted(os.listdir(tmp)) == sorted(f"{k}{ext}" for k in r for ext in (".npz", ".wav"))
    o = GO.generate(C.state_dict(), hp, **copy.deepcopy(req), **kw)
    assert list(r) == list(o)
    for k in r:
        for i, (x, y) in enumerate(zip(r[k], o[k])):
            assert x.shape == y.shape
            assert np.array_equal(x, y) if i < 2 else np.abs(x - y).max() <= 2e-5, (k, i)


@needs_reference
def test_symbol_table_is_the_references():
    R.import_reference_generate()
    from daft_exprt.symbols import symbols_english
    assert C.hparams().symbols == list(symbols_english)
    assert len(symbols_english) == C.hparams().n_symbols


# ---------------------------------------------------------------- CPU: product host side ------------------------------
@pytest.mark.parametrize("pt,seed,with_factors", [("add", 1, True), ("multiply", 2, True), ("add", 3, False)])
def test_collate_equals_oracle(pt, seed, with_factors):
    hp = C.hparams()
    req = synthetic.synthesis_request(hp, 9, 3, 25, seed, pt, with_factors=with_factors)
    mine = G.collate_tensors(*_collate_args(req, pt), external_prosody=req["external_prosody"])
    orac = GO.collate(*_collate_args(req, pt), external_prosody=req["external_prosody"])
    _same_collate(mine, orac)
    assert G.collate_tensors(*_collate_args(req, pt))[7] is None


def test_collate_tie_order_and_single_utterance():
    hp = C.hparams()
    sentences = [[["AA1", "B"], " ", ["K"], "~"], [["T", "S"], ",", ["IY0"], "~"], [["D"], "~"], [["M", "N"], " ", ["OW2"], "~"]]
    args = (sentences, [None] * 4, [None] * 4, [None] * 4, "multiply", [1, 0, 1, 0], ["a", "b", "c", "d"], hp)
    _same_collate(G.collate_tensors(*args), GO.collate(*args))
    one = G.collate_tensors(sentences[:1], [None], [[0.5] * 5], [None], "add", [1], ["x"], hp)
    assert one[0].tolist() == [[hp.symbols.index(s) for s in ("AA1", "B", " ", "K", "~")]]
    assert one[2].tolist() == [[0.5] * 5] and one[3].tolist() == [[0.0] * 5] and one[4].tolist() == [5]


@needs_reference
@pytest.mark.parametrize("pt,seed", [("add", 4), ("multiply", 5)])
def test_collate_equals_reference(pt, seed):
    ref_gen = R.import_reference_generate()
    hp = C.hparams()
    req = synthetic.synthesis_request(hp, 11, 3, 30, seed, pt)
    mine = G.collate_tensors(*_collate_args(req, pt), external_prosody=req["external_prosody"])
    theirs = ref_gen.collate_tensors(*_collate_args(req, pt), external_prosody=req["external_prosody"])
    _same_collate(mine, theirs)


PROSODY_CASES = [
    dict(pt="add", seed=6, src=None, alphas=(1.0, 1.0, 1.0)),
    dict(pt="add", seed=7, src=SRC, alphas=(1.0, 1.0, 1.0)),
    dict(pt="multiply", seed=8, src=SRC, alphas=(1.4, 0.6, 1.25)),
    dict(pt="multiply", seed=9, src=None, alphas=(0.5, 2.0, 0.1)),
]


def _prosody_inputs(case, n=10, speakers=None):
    hp = C.hparams()
    req = synthetic.synthesis_request(hp, n, 3, 28, case["seed"], case["pt"])
    if speakers is not None:
        req["speaker_ids"] = speakers
    col = GO.collate(*_collate_args(req, case["pt"]), external_prosody=req["external_prosody"])
    return hp, col


@pytest.mark.parametrize("case", PROSODY_CASES)
def test_external_prosody_tensors_equal_oracle(case):
    hp, col = _prosody_inputs(case)
    a_d, a_p, a_e = case["alphas"]
    mine = G.external_prosody_tensors(col[7], col[4], col[5], col[6], hp, col[0].shape[1], case["src"], a_d, a_p, a_e)
    orac = GO.external_tensors(col[7], col[4], col[5], hp, col[0].shape[1], case["src"], a_d, a_p, a_e)
    for k in ("duration_preds", "durations_int", "energy_preds", "pitch_preds"):
        assert mine[k].dtype == orac[k].dtype and torch.equal(mine[k], orac[k]), k


def test_external_prosody_speaker_fallback():
    """Unknown speakers use the statistics of 'spk 0' (generate.py:246-248)."""
    case = PROSODY_CASES[1]
    hp, col = _prosody_inputs(case, n=6, speakers=[0, 5, 1, 7, 0, 1])
    mine = G.external_prosody_tensors(col[7], col[4], col[5], col[6], hp, col[0].shape[1], case["src"])
    orac = GO.external_tensors(col[7], col[4], col[5], hp, col[0].shape[1], case["src"])
    for k in mine:
        assert torch.equal(mine[k], orac[k]), k


class _Spy:
    """Records what the reference driver hands to ``inference`` and forwards to the real model."""

    def __init__(self, model):
        self.model, self.calls = model, []

    def parameters(self):
        return self.model.parameters()

    def eval(self):
        return self.model.eval()

    def inference(self, inputs, pitch_transform, hparams, external_tensors, **kw):
        self.calls.append(([t.clone() for t in inputs], {k: v.clone() for k, v in external_tensors.items()}, dict(kw)))
        return self.model.inference(inputs, pitch_transform, hparams, external_tensors, **kw)


@needs_reference
@pytest.mark.parametrize("case", PROSODY_CASES)
def test_inference_arguments_equal_the_reference_drivers(case):
    """What reaches ``inference`` (inputs + external tensors) is bit-identical to what the reference's
    ``generate_batch_mel_specs`` builds (generate.py:199-284)."""
    hp = C.hparams()
    spy = _Spy(R.build_reference_model(hp, C.state_dict()))
    req = synthetic.synthesis_request(hp, 6, 3, 16, case["seed"], case["pt"])
    a_d, a_p, a_e = case["alphas"]
    with tempfile.TemporaryDirectory() as tmp:
        R.run_reference_generate(spy, hp, tmp, batch_size=6, source_stats=case["src"], alpha_dur=a_d, alpha_pitch=a_p,
                                 alpha_energy=a_e, **copy.deepcopy(req))
    (inputs, ext, kw), = spy.calls
    names = [f"{n}_spk_{s}" for n, s in zip(req["file_names"], req["speaker_ids"])]
    args = list(_collate_args(req, case["pt"]))
    args[6] = names
    col = G.collate_tensors(*args, external_prosody=req["external_prosody"])
    for k in range(6):
        assert col[k].dtype == inputs[k].dtype and torch.equal(col[k], inputs[k]), k
    mine = G.external_prosody_tensors(col[7], col[4], col[5], col[6], hp, col[0].shape[1], case["src"], a_d, a_p, a_e)
    for k in ext:
        assert mine[k].dtype == ext[k].dtype and torch.equal(mine[k], ext[k]), k
    # embeddings go through untouched: by position, not permuted by the length sort
    assert torch.equal(kw["external_embeddings"], req["external_embeddings"])
    assert torch.equal(kw["external_accent_emb"], req["external_accent_emb"])


@pytest.mark.parametrize("src", [None, {"mean": 1.1, "std": 0.7}])
def test_normalize_external_feature(src):
    g = torch.Generator().manual_seed(3)
    v = torch.randn(40, generator=g) * (torch.rand(40, generator=g) > 0.3)
    tgt = {"mean": 0.4, "std": 1.3}
    mine = G._normalize_external_feature(v, v == 0, tgt, src)
    assert torch.equal(mine, GO.normalize_feature(v, v == 0, tgt, src))
    assert torch.equal(mine == 0, v == 0) or src is not None
    if R.reference_available():
        assert torch.equal(mine, R.import_reference_generate()._normalize_external_feature(v, v == 0, tgt, src))
    with pytest.raises(ValueError, match="Target speaker stats std cannot be 0"):
        G._normalize_external_feature(v, v == 0, {"mean": 0.0, "std": 0}, src)
    with pytest.raises(ValueError, match="Source stats std cannot be 0"):
        G._normalize_external_feature(v, v == 0, tgt, {"mean": 0.0, "std": 0})


def test_driver_errors_follow_the_reference():
    hp = C.hparams()
    req = synthetic.synthesis_request(hp, 3, 4, 8, 12, "add")
    args = list(_collate_args(req, "add"))
    bad = copy.deepcopy(args)
    bad[0][1][0] = ["NOT_A_PHONE"]
    with pytest.raises(ValueError, match="is not in list"):
        G.collate_tensors(*bad)
    bad = copy.deepcopy(args)
    bad[1][0] = bad[1][0][:-1]
    with pytest.raises(AssertionError):
        G.collate_tensors(*bad)
    with pytest.raises(AssertionError):
        G.collate_tensors(*args, external_prosody=req["external_prosody"][:2])
    col = G.collate_tensors(*args, external_prosody=req["external_prosody"])
    L = col[0].shape[1]
    ok = (col[7], col[4], col[5], col[6], hp, L)
    short = copy.deepcopy(col[7])
    short[1]["symbols"] = short[1]["symbols"][:-1]
    with pytest.raises(AssertionError):
        G.external_prosody_tensors(short, *ok[1:])
    hp2 = C.hparams()
    hp2.stats = {"spk 3": hp2.stats["spk 0"]}
    with pytest.raises(KeyError, match="Speaker stats missing"):
        G.external_prosody_tensors(col[7], col[4], col[5], col[6], hp2, L)
    hp3 = C.hparams()
    hp3.stats["spk 0"]["pitch"]["std"] = 0
    hp3.stats["spk 1"]["pitch"]["std"] = 0
    with pytest.raises(ValueError, match="Speaker stats not initialized"):
        G.external_prosody_tensors(col[7], col[4], col[5], col[6], hp3, L)
    with pytest.raises(ValueError, match="Source stats std cannot be 0"):
        G.external_prosody_tensors(*ok, source_stats={"energy": {"mean": 0.0, "std": 0}, "pitch": {"mean": 0.0, "std": 1.0}})
    with pytest.raises(TypeError):          # generate_mel_specs' default alpha=None reaches the arithmetic (generate.py:231)
        G.external_prosody_tensors(*ok, alpha_dur=None)
    with pytest.raises(RuntimeError, match="runs the B200 kernels"):
        class _CpuModel:
            device = torch.device("cpu")
        G.launch_batch_mel_specs(_CpuModel(), *args[:7], hp)


def test_stage_block_layout():
    st = G._Stage(5, 13, 192, 128, pinned=False)
    offs = sorted((off, nbytes) for off, nbytes, _, _ in st.layout.values())
    assert all(off % 256 == 0 for off, _ in offs)
    assert all(a + n <= b for (a, n), (b, _) in zip(offs, offs[1:])) and offs[-1][0] + offs[-1][1] <= st.nbytes
    st.a["symbols"][:] = 7
    st.a["pitch_preds"][:] = 1.5
    mirror = st.block.clone()
    v = st.views(mirror)
    assert v["symbols"].dtype == torch.int64 and v["symbols"].shape == (5, 13) and int(v["symbols"].sum()) == 7 * 65
    assert v["pitch_preds"].dtype == torch.float32 and float(v["pitch_preds"].sum()) == 1.5 * 65
    assert v["external_embeddings"].shape == (5, 192) and v["external_accent_emb"].shape == (5, 128)


# ---------------------------------------------------------------- GPU: the product driver -----------------------------
def _model(mode):
    from daft_exprt_b200.model import DaftExprt
    m = DaftExprt(C.hparams())
    m.load_state_dict(C.state_dict())
    return m.cuda().eval().set_mode(mode)


def _check_predictions(preds, expect, mode):
    assert list(preds) == list(expect)
    num = den = 0.0
    for k in preds:
        p, e = preds[k], expect[k]
        for i in range(6):
            assert p[i].shape == e[i].shape and p[i].dtype == e[i].dtype, (k, i)
        assert np.array_equal(p[0], e[0]) and np.array_equal(p[1], e[1]), k        # durations: exact
        assert np.abs(p[2] - e[2]).max() <= 1e-5 and np.abs(p[3] - e[3]).max() <= 1e-4, k
        if mode == "fp32":
            assert np.abs(p[4] - e[4]).max() <= 1e-3, k                             # mel: north-star fp32 tolerance
            assert np.abs(p[5] - e[5]).max() <= 1e-4, k
        num += float(np.abs(p[4] - e[4]).sum())
        den += float(np.abs(e[4]).sum())
    assert num / den <= (2e-2 if mode == "bf16" else 1e-4)                           # bf16: relative L1 <= 2e-2


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["fp32", "bf16"])
@pytest.mark.parametrize("name", list(MGG.GENERATE_CASES))
def test_generate_mel_specs_equals_golden_and_oracle(name, mode):
    rk, gk = MGG.GENERATE_CASES[name]
    hp = C.hparams()
    req = synthetic.synthesis_request(hp, **rk)
    g = C.load_golden(name)
    assert str(g["request_digest"]) == MGG.request_digest(req)
    preds = G.generate_mel_specs(_model(mode), output_dir=None, hparams=hp, **copy.deepcopy(req), **gk)
    golden = {str(n): [g[f"{n}/{f}"] for f in MGG.FIELDS] for n in g["names"]}
    _check_predictions(preds, golden, mode)
    _check_predictions(preds, GO.generate(C.state_dict(), hp, **copy.deepcopy(req), **gk), mode)


@pytest.mark.gpu
def test_generate_files_vocoder_and_batch_call():
    hp = C.hparams()
    model = _model("fp32")
    req = synthetic.synthesis_request(hp, 5, 4, 12, 21, "add")
    kw = dict(batch_size=2, alpha_dur=1.0, alpha_pitch=1.0, alpha_energy=1.0)
    plain = G.generate_mel_specs(model, output_dir=None, hparams=hp, **copy.deepcopy(req), **kw)
    with tempfile.TemporaryDirectory() as tmp:
        with pytest.raises(ValueError, match="HiFi-GAN vocoder required"):          # generate.py:317-321
            G.generate_mel_specs(model, output_dir=tmp, hparams=hp, **copy.deepcopy(req), **kw)
        first = list(plain)[:2]
        assert sorted(os.listdir(tmp)) == sorted(f"{k}.npz" for k in first)      # written before the error, like there
    with tempfile.TemporaryDirectory() as tmp:
        preds = G.generate_mel_specs(model, output_dir=tmp, hparams=hp, vocoder=R._ZeroVocoder(), get_time_perf=True,
                                     **copy.deepcopy(req), **kw)
        assert sorted(os.listdir(tmp)) == sorted(f"{k}{ext}" for k in preds for ext in (".npz", ".wav"))
        for k in preds:
            assert np.array_equal(np.load(os.path.join(tmp, k + ".npz"))["mel_spec"], preds[k][4])
    # pipelined batches == one generate_batch_mel_specs call per batch, bit for bit
    for k in plain:
        assert all(np.array_equal(x, y) for x, y in zip(plain[k], preds[k]))
    names = list(req["file_names"][:2])
    with tempfile.TemporaryDirectory() as tmp:
        one = G.generate_batch_mel_specs(
            model, req["sentences"][:2], req["dur_factors"][:2], req["energy_factors"][:2], req["pitch_factors"][1][:2],
            "add", req["speaker_ids"][:2], names, tmp, hp, 1, req["external_prosody"][:2], R._ZeroVocoder(),
            external_embeddings=req["external_embeddings"][:2].cuda(), external_accent_emb=req["external_accent_emb"][:2])
    assert names == [f"{n}_spk_{s}" for n, s in zip(req["file_names"][:2], req["speaker_ids"][:2])]   # renamed in place
    for k in one:
        assert all(np.array_equal(x, y) for x, y in zip(one[k], plain[k]))


@pytest.mark.gpu
def test_generate_full_size_matches_direct_inference():
    """64 utterances of 20-150 symbols (BASELINE config 2's shape): the driver's crops equal a direct ``inference`` call on
    the tensors it builds; lengths and shapes are consistent."""
    hp = C.hparams()
    model = _model("bf16")
    req = synthetic.synthesis_request(hp, 64, 20, 150, 42, "add")
    kw = dict(alpha_dur=1.1, alpha_pitch=0.9, alpha_energy=1.05, source_stats=SRC)
    preds = G.generate_mel_specs(model, output_dir=None, hparams=hp, batch_size=64, **copy.deepcopy(req), **kw)
    names = [f"{n}_spk_{s}" for n, s in zip(req["file_names"], req["speaker_ids"])]
    col = G.collate_tensors(req["sentences"], req["dur_factors"], req["energy_factors"], req["pitch_factors"][1], "add",
                            req["speaker_ids"], names, hp, external_prosody=req["external_prosody"])
    ext = G.external_prosody_tensors(col[7], col[4], col[5], col[6], hp, col[0].shape[1], SRC, 1.1, 0.9, 1.05)
    enc, dec, w = model.inference(col[:6], "add", hp, ext, external_embeddings=req["external_embeddings"],
                                  external_accent_emb=req["external_accent_emb"])
    assert list(preds) == col[6] and len(preds) == 64
    out_len = dec[1].cpu()
    for i, k in enumerate(col[6]):
        n, t = int(col[4][i]), int(out_len[i])
        p = preds[k]
        assert p[4].shape == (hp.n_mel_channels, t) and p[5].shape == (n, t) and int(p[1].sum()) == t
        assert np.array_equal(p[1], enc[1][i, :n].cpu().numpy())
        assert np.array_equal(p[4], dec[0][i, :, :t].cpu().numpy())
        assert np.array_equal(p[5], w[i, :n, :t].cpu().numpy())
