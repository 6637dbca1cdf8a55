"""CPU: the writer half of the fine-tuning data set generation (fine_tune.py:77-128) against the
live reference loop: same files, byte for byte, same skip decisions and counters; partial-write clean-up; BulkRunner as
the producer."""
import copy
import os
import types

import numpy as np
import pytest
import torch

import _cases as C
from daft_exprt_b200 import fine_tune as FT
from oracle import ref_harness as RH

needs_reference = pytest.mark.skipif(not RH.reference_available(), reason="reference not present")


def _dataset(tmp_path, hp, spec):
    """spec: list of (speaker, name, seconds_total, begin_s, end_s).  Writes wavs + markers; returns items."""
    from scipy.io.wavfile import write
    rng = np.random.default_rng(5)
    items = []
    for speaker, name, secs, b, e in spec:
        os.makedirs(tmp_path / "data" / speaker / "wavs", exist_ok=True)
        fdir = tmp_path / "features" / speaker
        os.makedirs(fdir, exist_ok=True)
        wav = (rng.standard_normal(int(secs * hp.sampling_rate)) * 3000).astype(np.int16)
        write(str(tmp_path / "data" / speaker / "wavs" / f"{name}.wav"), hp.sampling_rate, wav)
        with open(fdir / f"{name}.markers", "w", encoding="utf-8") as f:
            mid = 0.5 * (b + e)
            f.write(f"{b}\t{mid}\tfoo\tw\n{mid}\t{e}\tbar\tw\n")
        items.append((str(fdir), name))
    return items


def _hp(tmp_path):
    hp = copy.deepcopy(C.hparams())
    hp.speakers = ["spk_a", "spk_b"]
    hp.data_set_dir = str(tmp_path / "data")
    hp.checkpoint = "x.pt"
    os.makedirs(tmp_path / "exp_ref", exist_ok=True)
    hp.training_files = str(tmp_path / "exp_ref" / "train.txt")
    for k, v in dict(mel_fmin=0.0, mel_fmax=8000.0, min_clipping=1e-5).items():
        if not hasattr(hp, k):
            setattr(hp, k, v)
    return hp


def _frames(hp, secs):
    return FT.hifi_mel_frames(int(int(secs[1] * hp.sampling_rate) - int(secs[0] * hp.sampling_rate)), hp)


@needs_reference
def test_writer_equals_the_reference_loop_byte_for_byte(tmp_path):
    hp = _hp(tmp_path)
    spec = [("spk_a", "ok_long", 3.0, 0.25, 2.5),        # written
            ("spk_b", "shape_off", 2.5, 0.1, 2.2),        # prediction one frame short -> skipped (shape)
            ("spk_a", "too_short", 1.5, 0.2, 0.9),        # 0.7 s -> skipped (< 1 s)
            ("spk_b", "ok_2", 2.0, 0.0, 1.6)]             # written
    items = _dataset(tmp_path, hp, spec)
    T = [_frames(hp, (b, e)) for _, _, _, b, e in spec]
    T[1] -= 1
    g = torch.Generator().manual_seed(3)
    preds = torch.randn(4, hp.n_mel_channels, max(T) + 7, generator=g)
    out_len = torch.tensor(T)
    batch = [None] * 11 + [[d for d, _ in items], [n for _, n in items]]
    # two loader batches, as the reference would see them
    batches = [tuple(batch[:11] + [batch[11][:2], batch[12][:2]]), tuple(batch[:11] + [batch[11][2:], batch[12][2:]])]
    predictions = [(preds[:2], out_len[:2]), (preds[2:], out_len[2:])]
    ref_dir = RH.run_reference_fine_tuning(hp, batches, predictions)
    ours_dir = str(tmp_path / "exp_ours" / "fine_tuning_dataset")
    w = FT.FineTuneWriter(hp, items, ft_data_set=ours_dir)
    for i in range(4):
        w(i, preds[i, :, :T[i]].numpy())
    assert (w.n_written, w.n_skipped_shape, w.n_skipped_short) == (2, 1, 1)
    ref_files = sorted(os.path.relpath(os.path.join(r, f), ref_dir) for r, _, fs in os.walk(ref_dir) for f in fs)
    our_files = sorted(os.path.relpath(os.path.join(r, f), ours_dir) for r, _, fs in os.walk(ours_dir) for f in fs)
    assert ref_files == our_files == ["spk_a/ok_long.npy", "spk_a/ok_long.wav", "spk_b/ok_2.npy", "spk_b/ok_2.wav"]
    for f in ref_files:
        assert open(os.path.join(ref_dir, f), "rb").read() == open(os.path.join(ours_dir, f), "rb").read(), f


def test_frame_count_formula_is_the_stft_of_the_reference_recipe():
    hp = C.hparams()
    n_fft, hop = int(hp.filter_length), int(hp.hop_length)
    pad = int((n_fft - hop) / 2)
    for n in (pad + 1, 1000, 22050, 22051, 48000, 65537):
        wav = torch.zeros(n)
        x = torch.nn.functional.pad(wav[None, None], (pad, pad), mode="reflect")[0, 0]
        spec = torch.stft(x, n_fft, hop_length=hop, win_length=n_fft, window=torch.hann_window(n_fft), center=False,
                          return_complex=True)
        assert spec.shape[1] == FT.hifi_mel_frames(n, hp), n
    with pytest.raises(ValueError):
        FT.hifi_mel_frames(pad, hp)


def test_partial_files_are_removed_and_the_run_goes_on(tmp_path, monkeypatch):
    hp = _hp(tmp_path)
    items = _dataset(tmp_path, hp, [("spk_a", "u0", 2.0, 0.0, 1.5), ("spk_a", "u1", 2.0, 0.0, 1.5)])
    T = _frames(hp, (0.0, 1.5))
    out = str(tmp_path / "ft")
    w = FT.FineTuneWriter(hp, items, ft_data_set=out)
    import scipy.io.wavfile as W
    real = W.write
    calls = []

    def failing(path, fs, x):
        calls.append(path)
        if len(calls) == 1:
            open(path, "wb").write(b"partial")
            raise OSError("disk full")
        return real(path, fs, x)

    monkeypatch.setattr(W, "write", failing)
    mel = np.zeros((hp.n_mel_channels, T), dtype=np.float32)
    w(0, mel)
    w(1, mel)
    assert w.n_written == 1
    assert sorted(os.listdir(os.path.join(out, "spk_a"))) == ["u1.npy", "u1.wav"]      # u0's .npy and partial .wav are gone


def test_unknown_or_ambiguous_speaker_is_the_reference_assertion(tmp_path):
    hp = _hp(tmp_path)
    w = FT.FineTuneWriter(hp, [(str(tmp_path / "features" / "nobody"), "x")], ft_data_set=str(tmp_path / "ft"))
    with pytest.raises(AssertionError):
        w(0, np.zeros((hp.n_mel_channels, 4), dtype=np.float32))


def test_bulk_runner_feeds_the_writer(tmp_path):
    """The runner's writer thread is the producer: every utterance of the shard reaches the sink with its cropped mel
    (CPU stand-in model, as in tests/test_bulk_sharding.py)."""
    from daft_exprt_b200 import bulk
    hp = _hp(tmp_path)
    n = 6
    stream = C.synthetic.UtteranceStream(hp, n, seed=11, lmin=4, lmax=9)
    in_lens, out_lens = stream.lengths()
    secs = [(t - 1) * hp.hop_length / hp.sampling_rate + 0.01 for t in out_lens]
    spec = [("spk_a" if i % 2 else "spk_b", f"u{i}", s + 0.2, 0.0, s) for i, s in enumerate(secs)]
    items = _dataset(tmp_path, hp, spec)

    class Echo(torch.nn.Module):
        device = torch.device("cpu")

        def forward(self, inputs):
            return None, None, None, (inputs[8] * 2.0, inputs[9]), None

    batches, mine = bulk.plan_shards(in_lens, out_lens, 4, 1)
    w = FT.FineTuneWriter(hp, items, ft_data_set=str(tmp_path / "ft"))
    frames = bulk.BulkRunner(Echo(), sink=w).run(stream, batches, mine[0])
    assert frames == sum(out_lens)
    long_enough = sum(1 for s in secs if int(s * hp.sampling_rate) >= hp.sampling_rate)
    assert w.n_written + w.n_skipped_shape + w.n_skipped_short == n and w.n_skipped_short == n - long_enough - w.n_skipped_shape
    for (spk, name, *_), t in zip(spec, out_lens):
        f = tmp_path / "ft" / spk / f"{name}.npy"
        if f.exists():
            assert np.load(f).shape == (hp.n_mel_channels, t)
