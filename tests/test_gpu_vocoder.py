"""GPU: the batched HiFi-GAN generator (dx_vocoder_forward through the Python mirror) against golden vectors frozen
from the live reference generator and against the CPU oracle.  Tolerances follow the north star's two modes:
fp32 mode max-abs <= 1e-3, bf16 mode relative L1 <= 2e-2 (waveform samples, tanh range)."""
import numpy as np
import pytest
import torch

import _cases as C
from oracle import hifigan_oracle as H
import make_vocoder_golden as VG
from daft_exprt_b200 import vocoder as V

pytestmark = pytest.mark.gpu
FP32_MAX_ABS = 1e-3
BF16_REL_L1 = 2e-2


def _generator(mode, seed=VG.WEIGHT_SEED):
    gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG)
    gen.load_state_dict(H.make_state_dict(seed=seed))
    return gen.cuda().eval().set_mode(mode)


def _check(audio, ref, mode):
    assert audio.shape == ref.shape
    assert np.isfinite(audio).all()
    if mode == "fp32":
        assert np.abs(audio - ref).max() <= FP32_MAX_ABS, np.abs(audio - ref).max()
    else:
        rel = np.abs(audio - ref).sum() / np.abs(ref).sum()
        assert rel <= BF16_REL_L1, rel


@pytest.mark.parametrize("mode", ["fp32", "bf16", "bf16+fp32res"])
@pytest.mark.parametrize("name", sorted(VG.VOCODER_CASES))
def test_generator_matches_golden(name, mode):
    """bf16 mode with its default bf16 residual stream and with the fp32 residual option (DX_VC_RESIDUAL_F32)."""
    g = C.load_golden(name)
    mels, lengths = VG.vocoder_case(**VG.VOCODER_CASES[name])
    gen = _generator(mode.split("+")[0]).set_residual("fp32" if mode.endswith("fp32res") else "bf16")
    mode = mode.split("+")[0]
    audio = gen(mels.cuda(), lengths)
    torch.cuda.synchronize()
    assert audio.shape == (mels.size(0), 1, mels.size(2) * 256)
    _check(audio[:, 0].cpu().numpy(), g["audio"], mode)


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_batch_equals_single_utterance_calls_bitwise(mode):
    """An utterance's samples do not depend on its neighbours in the packed stream, on its position, or on how the
    stream is tiled: the batched call and B single calls agree bit for bit, and so do two runs."""
    lengths = [40, 3, 31, 1, 18]
    mels, ln = VG.vocoder_case(lengths=lengths, seed=31)
    gen = _generator(mode)
    batch = gen(mels.cuda(), ln).cpu()
    again = gen(mels.cuda(), ln.cuda(), total_frames=sum(lengths)).cpu()
    assert torch.equal(batch, again)
    for b, n in enumerate(lengths):
        one = gen(mels[b : b + 1, :, :n].cuda()).cpu()          # lengths=None: the whole tensor, as the reference module
        assert torch.equal(one[0, 0], batch[b, 0, : n * 256]), b
        assert torch.count_nonzero(batch[b, 0, n * 256 :]) == 0


def test_oracle_parity_on_a_wider_batch_bf16_and_fp32():
    lengths = [120, 77, 64, 9]
    mels, ln = VG.vocoder_case(lengths=lengths, seed=33)
    sd = H.make_state_dict(seed=99)
    ref = H.infer_batch(sd, mels, ln).numpy()
    for mode in ("fp32", "bf16"):
        gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG)
        gen.load_state_dict(H.fold_weight_norm(sd))              # plain-weight checkpoint form
        gen.cuda().set_mode(mode)
        audio = gen(mels.cuda(), ln)[:, 0].cpu().numpy()
        _check(audio, ref, mode)


def test_vocoder_wrapper_contract():
    sd = H.make_state_dict(seed=VG.WEIGHT_SEED)
    voc = V.HiFiGanVocoder(state_dict=sd, device="cuda", mode="fp32")
    mels, ln = VG.vocoder_case(**VG.VOCODER_CASES["voc_ragged"])
    g = C.load_golden("voc_ragged")
    a = voc.infer(mels[1, :, :9].numpy())                        # numpy (80, T) in, clipped numpy out
    assert isinstance(a, np.ndarray) and a.shape == (9 * 256,) and np.abs(a).max() <= 1.0
    assert np.abs(a - g["audio"][1, : 9 * 256]).max() <= FP32_MAX_ABS
    with pytest.raises(ValueError):
        voc.infer(mels)                                          # batches are refused, as in the reference helper
    outs = voc.infer_batch(mels, ln)
    assert [o.shape[0] for o in outs] == [int(n) * 256 for n in ln]
    for b, o in enumerate(outs):
        assert np.abs(o - g["audio"][b, : o.shape[0]]).max() <= FP32_MAX_ABS


def test_synthesis_driver_writes_waves_through_the_batched_vocoder(tmp_path):
    """generate_mel_specs (row N1) with the B200 vocoder (row N4): one generator pass per batch; the .wav files equal
    the reference driver's per-utterance ``vocoder.infer`` conversion (generate.py:322-323) of the same mels."""
    import copy
    from scipy.io import wavfile
    from daft_exprt_b200 import generate as G, synthetic
    from daft_exprt_b200.model import DaftExprt
    hp = C.hparams()
    model = DaftExprt(hp).cuda()
    model.load_state_dict(C.state_dict())
    model.eval().set_mode("bf16")
    voc = V.HiFiGanVocoder(state_dict=H.make_state_dict(seed=VG.WEIGHT_SEED), device="cuda", mode="bf16")
    req = synthetic.synthesis_request(hp, 5, 4, 12, 21, "add")
    preds = G.generate_mel_specs(model, output_dir=str(tmp_path), hparams=hp, vocoder=voc, batch_size=3,
                                 alpha_dur=1.0, alpha_pitch=1.0, alpha_energy=1.0, **copy.deepcopy(req))
    assert len(preds) == 5
    for name, p in preds.items():
        sr, wav = wavfile.read(str(tmp_path / f"{name}.wav"))
        assert sr == hp.sampling_rate and wav.dtype == np.int16 and wav.shape[0] == p[4].shape[1] * 256
        audio = voc.infer(p[4])
        ref = (audio * 32767.5).clip(min=-32768, max=32767).astype(np.int16)
        assert np.array_equal(wav, ref), name


def test_fused_layers_equal_two_launch_layers_bitwise():
    """The fused ResBlock-layer kernel (hidden tile in shared memory) and the two-launch path (debug knob 12) run the
    same MMAs in the same order: identical bits."""
    from daft_exprt_b200 import _lib
    lib = _lib.load()
    lengths = [70, 33, 5]
    mels, ln = VG.vocoder_case(lengths=lengths, seed=41)
    gen = _generator("bf16")
    fused = gen(mels.cuda(), ln).cpu()
    try:
        lib.dx_set_debug(12, 1)
        plain = gen(mels.cuda(), ln).cpu()
    finally:
        lib.dx_set_debug(12, 0)
    assert torch.equal(fused, plain)
    gen.set_residual("fp32")                                    # same property with the fp32 residual stream
    fused32 = gen(mels.cuda(), ln).cpu()
    try:
        lib.dx_set_debug(12, 1)
        plain32 = gen(mels.cuda(), ln).cpu()
    finally:
        lib.dx_set_debug(12, 0)
    assert torch.equal(fused32, plain32) and not torch.equal(fused32, fused)


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
def test_empty_and_single_frame_utterances(mode):
    """Edge cases of the packed frame stream: a zero-length utterance produces silence and does not disturb its
    neighbours; 1-frame utterances (shorter than every receptive field) match the oracle."""
    lengths = [6, 0, 1, 1, 9]
    mels, ln = VG.vocoder_case(lengths=lengths, seed=43)
    sd = H.make_state_dict(seed=VG.WEIGHT_SEED)
    gen = _generator(mode)
    audio = gen(mels.cuda(), ln)[:, 0].cpu()
    assert torch.count_nonzero(audio[1]) == 0
    ref = H.infer_batch(sd, mels, ln)
    _check(audio.numpy(), ref.numpy(), mode)
    alone = gen(mels[4:5, :, :9].cuda())[0, 0].cpu()
    assert torch.equal(alone, audio[4, : 9 * 256])


def test_long_form_utterance_properties():
    """cfg4-sized input (2000 frames = 512,000 samples, 23 s of audio) beside a short one: finite, in tanh range,
    deterministic, identical to the single call, and the first 64 frames' samples (receptive field of the generator
    is < 32 frames) match the oracle run on a 96-frame prefix."""
    lengths = [2000, 150]
    mels, ln = VG.vocoder_case(lengths=lengths, seed=47)
    gen = _generator("bf16")
    a = gen(mels.cuda(), ln)[:, 0]
    b = gen(mels.cuda(), ln)[:, 0]
    assert torch.equal(a, b) and torch.isfinite(a).all() and float(a.abs().max()) <= 1.0
    one = gen(mels[1:2, :, :150].cuda())[0, 0]
    assert torch.equal(one, a[1, : 150 * 256])
    sd = H.make_state_dict(seed=VG.WEIGHT_SEED)
    ref = H.infer(sd, mels[0, :, :96])[: 64 * 256].numpy()
    got = a[0, : 64 * 256].cpu().numpy()
    assert np.abs(got - ref).sum() / np.abs(ref).sum() <= BF16_REL_L1


def test_multi_pass_batches_equal_one_pass():
    """Batches above max_frames_per_pass are vocoded group by group: same bits as one pass."""
    lengths = [30, 28, 20, 9, 3]
    mels, ln = VG.vocoder_case(lengths=lengths, seed=51)
    gen = _generator("bf16")
    one = gen(mels.cuda(), ln).cpu()
    gen.max_frames_per_pass = 50
    many = gen(mels.cuda(), ln).cpu()
    assert torch.equal(one, many)


@pytest.mark.parametrize("seed", [101, 102, 103])
def test_random_batch_compositions(seed):
    """Property sweep: random batch sizes and lengths (including empty utterances and tile-boundary lengths around
    118 / 122 / 126 = the fused kernel's valid rows per tile at 8x ... 256x the frame rate): fp32 mode against the
    oracle, bf16 mode batched == single calls bitwise."""
    g = torch.Generator().manual_seed(seed)
    B = int(torch.randint(2, 9, (1,), generator=g))
    lengths = [int(v) for v in torch.randint(0, 40, (B,), generator=g)]
    lengths[0] = max(lengths[0], 1)
    lengths[-1] = [15, 29, 59][seed % 3]            # 59 frames * 2 = 118 rows at the first fused stage's rate ... etc.
    mels, ln = VG.vocoder_case(lengths=lengths, seed=seed)
    sd = H.make_state_dict(seed=VG.WEIGHT_SEED)
    ref = H.infer_batch(sd, mels, ln).numpy()
    _check(_generator("fp32")(mels.cuda(), ln)[:, 0].cpu().numpy(), ref, "fp32")
    gen = _generator("bf16")
    batch = gen(mels.cuda(), ln).cpu()
    _check(batch[:, 0].numpy(), ref, "bf16")
    for b, n in enumerate(lengths):
        if n > 0:
            one = gen(mels[b : b + 1, :, :n].cuda()).cpu()
            assert torch.equal(one[0, 0], batch[b, 0, : n * 256]), (b, n)
        assert torch.count_nonzero(batch[b, 0, n * 256 :]) == 0
