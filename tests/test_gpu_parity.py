"""GPU parity tests proper: the CUDA path (through the reference-shaped Python mirror → C ABI) against
(1) the golden vectors frozen from the live reference and (2) the oracle port on the same seeded inputs.

Gates (BASELINE.json north_star): integer length-regulator quantities bit-exact; fp32 mode mel max-abs <= 1e-3;
bf16 mode relative L1 <= 2e-2 (over valid frames; padded frames must be exactly 0)."""
import pytest
import torch

import _cases as C
from daft_exprt_b200.model import DaftExprt
from oracle import daft_oracle as O

pytestmark = pytest.mark.gpu

FP32_MAX_ABS = 1e-3
BF16_REL_L1 = 2e-2
# bf16 tolerances of the SECONDARY outputs (the north-star gate is the mel's relative L1 above): 2 x the worst value
# measured on B200 over every case of this file and test_gpu_round2.py (gpurun_out/parity_measured.json, copied to
# profiles/r02_parity_measured.json)
# measured worst -> tolerance: film_dec 2.7e-3 -> 6e-3; speaker_preds 2.5e-3 -> 6e-3; enc_out 3.8e-2 -> 8e-2; accent_emb 1.0e-2
# -> 2.1e-2; alignment weights: mean-abs 5.5e-4 -> 1.1e-3, 99.9th percentile 1.0e-2 -> 2e-2; the single worst element of the
# ~7 M valid ones is 0.28 (cfg2-size inference; a ratio of two Gaussians where the predicted range sigma is smallest, so
# the bf16 encoder output moves it most) -> 0.35 on the maximum (1.25 x: the maximum is one element, the percentile and the
# mean carry the gate)
W_BF16_MAX_ABS = 0.35
W_BF16_P999_ABS = 2e-2
W_BF16_MEAN_ABS = 1.1e-3
FILM_BF16_MAX_ABS = 6e-3
ENC_BF16_MAX_ABS = 8e-2
ACCENT_BF16_MAX_ABS = 2.1e-2


@pytest.fixture(scope="module")
def models():
    hp = C.hparams()
    out = {}
    for mode in ("fp32", "bf16"):
        m = DaftExprt(hp, mode=mode).cuda()
        m.load_state_dict(C.state_dict())
        out[mode] = m
    return out


def _valid_mask(lengths, width):
    return torch.arange(width).unsqueeze(0) < lengths.unsqueeze(1)


def _check_mel(mel, ref_mel, out_len, mode):
    mel = mel.float().cpu()
    valid = _valid_mask(out_len.cpu(), mel.shape[2]).unsqueeze(1).expand_as(mel)
    assert torch.all(mel[~valid] == 0), "padded frames must be exactly zero (FrameDecoder mask, model.py:562)"
    diff = (mel - ref_mel).abs()
    if mode == "fp32":
        C.record("mel_max_abs_fp32", diff[valid].max())
        assert diff[valid].max().item() <= FP32_MAX_ABS, f"fp32 mel max-abs {diff[valid].max().item():.3e}"
    else:
        rel = (diff[valid].sum() / ref_mel[valid].abs().sum()).item()
        C.record("mel_rel_l1_bf16", rel)
        assert rel <= BF16_REL_L1, f"bf16 mel rel-L1 {rel:.3e}"


def _check_weights(w, ref_w, in_len, out_len, mode):
    w = w.cpu()
    B, L, T = ref_w.shape
    lv = _valid_mask(in_len.cpu(), L).unsqueeze(2)
    tv = _valid_mask(out_len.cpu(), T).unsqueeze(1)
    valid = (lv & tv).expand_as(ref_w)
    # the bf16 gate of the north star is on the mel; the alignment is a softmax-like ratio that is steep where the
    # predicted range σ is small, so bf16 encoder outputs move it by several 1e-2 in places
    tol = 1e-4 if mode == "fp32" else W_BF16_MAX_ABS
    C.record(f"weights_max_abs_{mode}", (w - ref_w)[valid].abs().max())
    C.record(f"weights_mean_abs_{mode}", (w - ref_w)[valid].abs().mean())
    assert (w - ref_w)[valid].abs().max().item() <= tol
    if mode == "bf16":
        assert (w - ref_w)[valid].abs().mean().item() <= W_BF16_MEAN_ABS
        d = (w - ref_w)[valid].abs().flatten()
        p999 = d.kthvalue(max(1, int(0.999 * d.numel()))).values.item()
        C.record("weights_p999_abs_bf16", p999)
        assert p999 <= W_BF16_P999_ABS
    # padded symbols carry exactly zero weight (model.py:503)
    assert torch.all(w[(~lv).expand_as(w)] == 0)
    # column sums agree with the reference's (1 where any Gaussian reaches the frame; the reference's columns
    # underflow to 0 where every p[l, t] underflows, model.py:505)
    colsum, ref_colsum = (w * lv).sum(dim=1), (ref_w * lv).sum(dim=1)
    C.record(f"weights_colsum_max_abs_{mode}", (colsum - ref_colsum)[tv.squeeze(1)].abs().max())
    assert (colsum - ref_colsum)[tv.squeeze(1)].abs().max().item() <= (1e-4 if mode == "fp32" else W_BF16_MAX_ABS)


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
@pytest.mark.parametrize("name", sorted(C.MG.FORWARD_CASES))
def test_forward_vs_golden(models, mode, name):
    hp, batch, g = C.forward_case(name)
    m = models[mode]
    dev = tuple(t.cuda() for t in batch)
    spk, film, enc, dec, w = m(dev)
    torch.cuda.synchronize()
    assert torch.equal(dec[1].cpu(), C.t(g["output_lengths"]))
    _check_mel(dec[0], C.t(g["mel"]), batch[9], mode)
    _check_weights(w, C.t(g["weights"]), batch[5], batch[9], mode)
    tol = 1e-3 if mode == "fp32" else FILM_BF16_MAX_ABS
    C.record(f"film_dec_max_abs_{mode}", (film[3].cpu() - C.t(g["film_dec"])).abs().max())
    C.record(f"speaker_preds_max_abs_{mode}", (spk.cpu() - C.t(g["speaker_preds"])).abs().max())
    assert (film[3].cpu() - C.t(g["film_dec"])).abs().max().item() <= tol
    assert (spk.cpu() - C.t(g["speaker_preds"])).abs().max().item() <= tol
    # the tuple carries the caller's own tensors through, like the reference (model.py:943-946)
    assert enc[0] is dev[1] and enc[3] is dev[5] and dec[1] is dev[9]
    acc = m.accent_encoder(dev[6], dev[7], dev[8], dev[9]).cpu()
    C.record(f"accent_emb_max_abs_{mode}", (acc - C.t(g["accent_emb"])).abs().max())
    assert (acc - C.t(g["accent_emb"])).abs().max().item() <= (1e-4 if mode == "fp32" else ACCENT_BF16_MAX_ABS)


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
@pytest.mark.parametrize("name", sorted(C.MG.INFERENCE_CASES))
def test_inference_vs_golden(models, mode, name):
    hp, pt, inputs, prosody, ext_emb, acc_emb, g = C.inference_case(name)
    m = models[mode]
    dev_in = tuple(t.cuda() for t in inputs)
    dev_pros = {k: v.clone().cuda() for k, v in prosody.items()}
    enc, dec, w = m.inference(dev_in, pt, hp, dev_pros, ext_emb.cuda(), acc_emb.cuda())
    torch.cuda.synchronize()
    # integer contract: bit-exact in both modes
    assert torch.equal(enc[1].cpu(), C.t(g["durations_int"]))
    assert torch.equal(dec[1].cpu(), C.t(g["output_lengths"]))
    assert torch.equal(enc[0].cpu(), C.t(g["duration_preds"]))
    assert (enc[2].cpu() - C.t(g["energy_preds"])).abs().max().item() <= 1e-6
    assert (enc[3].cpu() - C.t(g["pitch_preds"])).abs().max().item() <= 2e-5
    # H8: the reference mutates the caller's pitch tensor
    assert torch.equal(dev_pros["pitch_preds"], enc[3])
    _check_mel(dec[0], C.t(g["mel"]), C.t(g["output_lengths"]), mode)
    _check_weights(w, C.t(g["weights"]), inputs[4], C.t(g["output_lengths"]), mode)


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
@pytest.mark.parametrize("kw", [
    dict(batch=1, lmin=1, lmax=1, seed=31, const_dur=1),            # single symbol, single frame
    dict(batch=2, lmin=2, lmax=3, seed=32),                         # tiny ragged
    dict(batch=7, lmin=5, lmax=61, seed=33),                        # ragged, crosses 128-row tiles
    dict(batch=3, lmin=40, lmax=40, seed=34, const_dur=4),          # no padding at all (true zero pad at the end)
    dict(batch=16, lmin=20, lmax=90, seed=35),                      # several tiles per utterance
])
def test_forward_vs_oracle(models, mode, kw):
    hp = C.hparams()
    batch = C.synthetic.forward_batch(hp, **kw)
    ref = O.forward(C.state_dict(), hp, tuple(t.clone() for t in batch))
    m = models[mode]
    out = m(tuple(t.cuda() for t in batch), _want_enc_out=True)
    torch.cuda.synchronize()
    _check_mel(out[3][0], ref[3][0], batch[9], mode)
    _check_weights(out[4], ref[4], batch[5], batch[9], mode)
    sd = C.state_dict()
    accent = O.accent_encoder(sd, hp, batch[6], batch[7], batch[8], batch[9])
    film = O.style_adapter(sd, hp, accent + O.speaker_projection(sd, batch[11]))
    enc_ref = O.phoneme_encoder(sd, hp, batch[0], film["phoneme_encoder"], batch[5])
    enc = m._last_debug["enc_out"].cpu()
    tol = 1e-3 if mode == "fp32" else ENC_BF16_MAX_ABS
    C.record(f"enc_out_max_abs_{mode}", (enc - enc_ref).abs().max())
    assert (enc - enc_ref).abs().max().item() <= tol


def test_forward_external_embeddings_and_errors(models):
    hp = C.hparams()
    m = models["fp32"]
    batch = C.synthetic.forward_batch(hp, batch=3, lmin=6, lmax=14, seed=41)
    g = torch.Generator().manual_seed(1)
    acc = torch.randn(3, 128, generator=g)
    spk = torch.randn(3, 128, generator=g)
    ref = O.forward(C.state_dict(), hp, tuple(t.clone() for t in batch), external_accent_emb=acc, external_spk_emb=spk)
    out = m(tuple(t.cuda() for t in batch), external_accent_emb=acc.cuda(), external_spk_emb=spk.cuda())
    _check_mel(out[3][0], ref[3][0], batch[9], "fp32")
    assert (out[0].cpu() - ref[0]).abs().max().item() <= 1e-4
    with pytest.raises(ValueError):
        m(tuple(t.cuda() for t in batch)[:11])
    no_spk = tuple(t.cuda() for t in batch[:11]) + (None,)
    with pytest.raises(ValueError):
        m(no_spk)
    inputs, prosody, ee, ae = C.synthetic.inference_batch(hp, 2, 5, 6, 1)
    dev_in = tuple(t.cuda() for t in inputs)
    pros = {k: v.cuda() for k, v in prosody.items()}
    with pytest.raises(ValueError):
        m.inference(dev_in, "add", hp, pros, None, ae.cuda())
    with pytest.raises(ValueError):
        m.inference(dev_in, "add", hp, pros, ee.cuda(), None)
    with pytest.raises(ValueError):
        m.inference(dev_in, "add", hp, None, ee.cuda(), ae.cuda())
    with pytest.raises(NotImplementedError):
        m.inference(dev_in, "bogus", hp, pros, ee.cuda(), ae.cuda())


def test_batch_composition_matters_like_the_reference(models):
    """H1: an utterance's result depends on whether it is padded in its batch.  Solo vs padded-batch results differ
    in the reference; ours must reproduce each, not an idealised padding-neutral result."""
    hp = C.hparams()
    m = models["fp32"]
    batch = C.synthetic.forward_batch(hp, batch=2, lmin=10, lmax=20, seed=52)
    if int(batch[5][0]) == int(batch[5][1]):
        pytest.skip("lengths equal")
    ref_b = O.forward(C.state_dict(), hp, tuple(t.clone() for t in batch))[3][0]
    # second (shorter) utterance alone, cropped to its own length
    l1, t1 = int(batch[5][1]), int(batch[9][1])
    solo = tuple([batch[0][1:2, :l1], batch[1][1:2, :l1], batch[2][1:2, :l1], batch[3][1:2, :l1], batch[4][1:2, :l1],
                  batch[5][1:2], batch[6][1:2, :t1], batch[7][1:2, :t1], batch[8][1:2, :, :t1], batch[9][1:2],
                  batch[10][1:2], batch[11][1:2]])
    ref_s = O.forward(C.state_dict(), hp, tuple(t.clone() for t in solo))[3][0]
    out_b = m(tuple(t.cuda() for t in batch))[3][0].cpu()
    out_s = m(tuple(t.contiguous().cuda() for t in solo))[3][0].cpu()
    assert (ref_b[1, :, :t1] - ref_s[0]).abs().max() > 1e-3          # the reference itself differs
    assert (out_b[1, :, :t1] - ref_b[1, :, :t1]).abs().max() <= FP32_MAX_ABS
    assert (out_s[0] - ref_s[0]).abs().max() <= FP32_MAX_ABS


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_cfg2_full_size_properties(models, mode):
    """BASELINE cfg2 (B=64, L in 20..150): size-independent properties + parity against the oracle."""
    hp = C.hparams()
    batch = C.synthetic.config_batch(hp, "cfg2")
    m = models[mode]
    out = m(tuple(t.cuda() for t in batch))
    torch.cuda.synchronize()
    mel, w = out[3][0].cpu(), out[4].cpu()
    assert torch.isfinite(mel).all() and torch.isfinite(w).all()
    out_len, in_len = batch[9], batch[5]
    valid = _valid_mask(out_len, mel.shape[2]).unsqueeze(1).expand_as(mel)
    assert torch.all(mel[~valid] == 0)
    assert w.min() >= 0 and w.max() <= 1 + 1e-5
    assert (w.sum(dim=1) <= 1 + 1e-4).all()
    # idempotence: same inputs, same result (no stale-workspace dependence)
    out2 = m(tuple(t.cuda() for t in batch))
    # both modes are run-to-run deterministic (no atomics anywhere on the path)
    assert torch.equal(out2[3][0].cpu(), mel)
    assert torch.equal(out2[4].cpu(), w)
    ref = O.forward(C.state_dict(), hp, tuple(t.clone() for t in batch))
    _check_mel(out[3][0], ref[3][0], out_len, mode)
    _check_weights(out[4], ref[4], in_len, out_len, mode)


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_cfg4_long_form(models, mode):
    """BASELINE config 4: 400 symbols, 2000 frames (attention-dominated regime; 16 KV tiles per query tile)."""
    hp = C.hparams()
    batch = C.synthetic.config_batch(hp, "cfg4_b1")
    assert batch[8].shape[2] == 2000
    ref = O.forward(C.state_dict(), hp, tuple(t.clone() for t in batch))
    out = models[mode](tuple(t.cuda() for t in batch))
    torch.cuda.synchronize()
    _check_mel(out[3][0], ref[3][0], batch[9], mode)
    _check_weights(out[4], ref[4], batch[5], batch[9], mode)


def test_bulk_runner_matches_direct_forward(models):
    """Config 3 plumbing on the device: pipelined H2D / forward / D2H + per-utterance crop == a direct call."""
    from daft_exprt_b200 import bulk
    hp = C.hparams()
    m = models["bf16"]
    stream = C.synthetic.UtteranceStream(hp, 24, seed=5, lmin=5, lmax=40)
    in_lens, out_lens = stream.lengths()
    batches, mine = bulk.plan_shards(in_lens, out_lens, 8, 1)
    got = {}
    frames = bulk.BulkRunner(m, sink=lambda i, mel: got.__setitem__(i, torch.from_numpy(mel.copy()))).run(stream, batches, mine[0])
    assert frames == sum(out_lens) and sorted(got) == list(range(24))
    for b in batches:
        host = bulk.collate(stream, b)
        mel = m(tuple(t.cuda() for t in host))[3][0].cpu()
        for k, i in enumerate(b):
            assert torch.equal(got[i], mel[k, :, :out_lens[i]])


@pytest.mark.parametrize("cfg", ["cfg2", "cfg4_b1"])
def test_fused_block_tail_matches_four_launch_blocks(models, cfg):
    """The fused block kernel (out-proj + LN1 + FFN + next QKV, two launches per FFT block) computes bit for bit what
    the four-launch block (debug knob 9: QKV, attention, out-proj + LN1, FFN) computes: same bf16 rounding points,
    same accumulation order.  cfg2 exercises two-half tiles (frames) and one-half tiles (symbols)."""
    from daft_exprt_b200 import _lib
    hp = C.hparams()
    batch = tuple(t.cuda() for t in C.synthetic.config_batch(hp, cfg))
    m = models["bf16"]
    lib = _lib.load()
    fused = m(batch)
    torch.cuda.synchronize()
    lib.dx_set_debug(9, 1)
    try:
        plain = m(batch)
        torch.cuda.synchronize()
    finally:
        lib.dx_set_debug(9, 0)
    assert torch.equal(fused[3][0], plain[3][0])          # mel
    assert torch.equal(fused[4], plain[4])                # gaussian upsampling weights
    assert torch.equal(fused[0], plain[0])                # accent embedding


@pytest.mark.parametrize("cfg", ["cfg1", "cfg2", "cfg4_b1"])
def test_hidden_dimension_split_of_the_fused_block_kernel(models, cfg):
    """The fused block kernel splits the 1024-wide hidden dimension of a tile across CTAs (stream-K; partial fp32 sums
    meet in the tile's owner in a fixed order).  Against whole tiles per CTA (debug knob 6) only the fp32 summation
    order of conv2 changes: the mel agrees to bf16 rounding, both stay inside the north-star gate against the oracle,
    and the split path is deterministic run to run.  The split is on by default for short streams only; knob 8 = 1000
    forces it for every stream (cfg2 frames: segments that span two tiles), knob 7 = 1 the deepest split (8 CTAs per
    tile)."""
    from daft_exprt_b200 import _lib
    hp = C.hparams()
    host = C.synthetic.config_batch(hp, cfg)
    batch = tuple(t.cuda() for t in host)
    ref = O.forward(C.state_dict(), hp, tuple(t.clone() for t in host)) if cfg != "cfg2" else None
    m = models["bf16"]
    lib = _lib.load()
    split = m(batch)
    split = (split[3][0].clone(), split[4].clone(), split[0].clone())
    again = m(batch)
    torch.cuda.synchronize()
    assert torch.equal(split[0], again[3][0]) and torch.equal(split[1], again[4]) and torch.equal(split[2], again[0])
    lib.dx_set_debug(6, 1)
    try:
        whole = m(batch)
        whole = (whole[3][0].clone(), whole[4].clone())
        torch.cuda.synchronize()
    finally:
        lib.dx_set_debug(6, 0)
    lib.dx_set_debug(8, 1000)
    try:
        forced = m(batch)
        forced = (forced[3][0].clone(), forced[4].clone())
        lib.dx_set_debug(7, 1)
        deep = m(batch)
        deep = (deep[3][0].clone(), deep[4].clone())
        torch.cuda.synchronize()
    finally:
        lib.dx_set_debug(7, 0)
        lib.dx_set_debug(8, 0)
    for other in (whole, forced, deep):
        rel = ((split[0] - other[0]).abs().sum() / other[0].abs().sum()).item()
        assert rel <= 1e-2, rel
    if ref is not None:
        for got in (split, whole, forced, deep):
            _check_mel(got[0], ref[3][0], host[9], "bf16")


def test_hidden_dimension_split_sweep_over_row_counts(models):
    """The split of the fused block kernel is the default for short streams (the serving case), and its segment walk has
    many boundary cases (a tile's chunks on 2..8 CTAs, segments that span two tiles, last tiles with a handful of rows,
    1-row streams).  Sweep the row count of both streams through those cases and compare with whole tiles per CTA (debug
    knob 6, itself checked against the oracle elsewhere): same outputs to bf16 rounding, nothing non-finite, exact zeros
    in the padding."""
    from daft_exprt_b200 import _lib
    hp = C.hparams()
    m = models["bf16"]
    lib = _lib.load()
    cases = [(1, 1, 1), (1, 2, 3), (1, 17, 7), (1, 60, 2), (1, 61, 2), (1, 62, 2), (1, 40, 3), (1, 41, 3), (1, 124, 1),
             (1, 125, 1), (2, 121, 1), (2, 122, 2), (3, 83, 3), (1, 200, 5), (1, 249, 1), (4, 30, 9), (5, 150, 1), (7, 9, 1),
             (2, 248, 2), (6, 100, 4)]
    worst = 0.0
    for knob7 in (0, 1):                   # default policy, then the deepest split
        for i, (B, L, d) in enumerate(cases):
            host = C.synthetic.forward_batch(hp, batch=B, lmin=max(1, L - 3 * (i % 3)), lmax=L, seed=300 + i, const_dur=d)
            batch = tuple(t.cuda() for t in host)
            lib.dx_set_debug(7, knob7)
            try:
                got = m(batch)
                got = (got[3][0].clone(), got[4].clone(), got[0].clone())
                lib.dx_set_debug(6, 1)
                want = m(batch)
                want = (want[3][0].clone(), want[4].clone(), want[0].clone())
                torch.cuda.synchronize()
            finally:
                lib.dx_set_debug(6, 0)
                lib.dx_set_debug(7, 0)
            for a, b in zip(got, want):
                assert torch.isfinite(a).all(), (B, L, d)
            rel = ((got[0] - want[0]).abs().sum() / want[0].abs().sum().clamp_min(1e-9)).item()
            worst = max(worst, rel)
            assert rel <= 2e-2, (B, L, d, knob7, rel)
            out_len = host[9]
            for b in range(B):
                assert float(got[0][b, :, int(out_len[b]):].abs().sum()) == 0.0, (B, L, d, "padding must be exactly zero")
    C.record("ffn_split_sweep_rel_l1_vs_whole_tiles", worst)
