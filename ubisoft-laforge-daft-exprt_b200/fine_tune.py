"""The writer half of the reference's vocoder fine-tuning data set generation.

Replaces the per-utterance body of ``fine_tuning``'s main loop, ``/root/reference/src/daft_exprt/fine_tune.py:77-128``:
crop of the predicted mel, the utterance's audio cut to its sentence markers, the target / prediction shape check, the
"at least one second" rule, ``np.save`` + ``scipy.io.wavfile.write`` and the removal of half-written files.  A
``FineTuneWriter`` is a ``sink`` for ``bulk.BulkRunner``: the runner's writer thread calls it with the cropped
``(80, T_b)`` prediction of one utterance while the GPU works on later batches, so file I/O is off the GPU's critical
path (the reference does all of it between two forwards, after a blocking ``.cpu()``).

Differences that do not change what is written:
  * the reference computes the target mel-spectrogram (``mel_spectrogram_HiFi``, ``extract_features.py:345-379``) only to
    compare its SHAPE with the prediction's; the frame count of that STFT (reflect padding ``(n_fft - hop) / 2`` per side,
    ``center=False``) is ``1 + (len + 2 pad - n_fft) // hop``, so the check is made on that number and no STFT runs;
  * audio is read by a pluggable ``wav_loader(path, sr) -> (float32 mono samples, sr)``; the default uses
    ``scipy.io.wavfile`` (``librosa`` is not a dependency of this package) and refuses files whose rate differs from
    ``hparams.sampling_rate`` instead of resampling them - pass ``librosa.load`` to get the reference's behaviour.
"""
import logging
import os
import threading

import numpy as np

_logger = logging.getLogger(__name__)


def default_wav_loader(path, sr):
    """``librosa.load(path, sr=sr)`` for files that are already at ``sr``: mono float32 in [-1, 1)."""
    from scipy.io import wavfile
    fs, x = wavfile.read(path)
    if fs != sr:
        raise ValueError(f"{path}: sampling rate {fs} != {sr} (resampling needs librosa: pass wav_loader=librosa.load)")
    if x.dtype == np.int16:
        y = x.astype(np.float32) / 32768.0
    elif x.dtype == np.int32:
        y = x.astype(np.float32) / 2147483648.0
    elif x.dtype == np.uint8:
        y = (x.astype(np.float32) - 128.0) / 128.0
    elif x.dtype in (np.float32, np.float64):
        y = x.astype(np.float32)
    else:
        raise TypeError(f"{path}: unsupported sample type {x.dtype}")
    if y.ndim == 2:                      # librosa's mono=True: mean of the channels
        y = y.mean(axis=1)
    return y, fs


def rescale_wav_to_float32(x):
    """``extract_features.py:382-404``."""
    if x.dtype == "int16":
        y = x / 32768.0
    elif x.dtype == "int32":
        y = x / 2147483648.0
    elif x.dtype == "uint8":
        y = ((x / 255.0) - 0.5) * 2
    elif x.dtype == "float32" or x.dtype == "float64":
        y = x
    else:
        raise TypeError(f"could not normalize wav, unsupported sample type {x.dtype}")
    return y.astype("float32")


def hifi_mel_frames(n_samples, hparams):
    """Frame count of ``mel_spectrogram_HiFi`` (``extract_features.py:361-371``) for ``n_samples`` samples."""
    n_fft, hop = int(hparams.filter_length), int(hparams.hop_length)
    pad = int((n_fft - hop) / 2)
    if n_samples <= pad:                 # torch's reflect padding needs pad < length: the reference raises here
        raise ValueError(f"audio of {n_samples} samples is shorter than the reflect padding ({pad})")
    return 1 + (n_samples + 2 * pad - n_fft) // hop


def read_markers_span(markers_file):
    """Sentence begin / end in seconds: first field of the first line, second field of the last (``fine_tune.py:96-100``)."""
    with open(markers_file, "r", encoding="utf-8") as f:
        lines = f.readlines()
    begin = float(lines[0].strip().split(sep="\t")[0])
    end = float(lines[-1].strip().split(sep="\t")[1])
    return begin, end


class FineTuneWriter:
    """``sink(utt_id, mel_pred (n_mels, T))`` that writes ``<ft_data_set>/<speaker>/<file>.npy`` and ``.wav``.

    ``items[utt_id] = (feature_dir, feature_file)`` - what the reference's collate returns at batch indices 11 and 12.
    Counters ``n_written`` / ``n_skipped_shape`` / ``n_skipped_short`` are the reference's summary line."""

    def __init__(self, hparams, items, ft_data_set=None, wav_loader=None):
        self.hp = hparams
        self.items = items
        if ft_data_set is None:         # fine_tune.py:53-55
            ft_data_set = os.path.join(os.path.dirname(hparams.training_files), "fine_tuning_dataset")
        self.ft_data_set = ft_data_set
        self.wav_loader = wav_loader or default_wav_loader
        self.n_written = self.n_skipped_shape = self.n_skipped_short = 0
        self._lock = threading.Lock()
        for speaker in hparams.speakers:            # fine_tune.py:56-57
            os.makedirs(os.path.join(ft_data_set, speaker), exist_ok=True)

    def __call__(self, utt_id, mel_spec_pred):
        hp = self.hp
        feature_dir, feature_file = self.items[utt_id]
        speaker_name = [speaker for speaker in hp.speakers if feature_dir.endswith(speaker)]
        assert len(speaker_name) == 1, f"{feature_dir} -- {feature_file} -- {speaker_name}"
        speaker_name = speaker_name[0]
        wav_file = os.path.join(hp.data_set_dir, speaker_name, "wavs", f"{feature_file}.wav")
        wav, fs = self.wav_loader(wav_file, hp.sampling_rate)
        wav = rescale_wav_to_float32(np.asarray(wav))
        sent_begin, sent_end = read_markers_span(os.path.join(feature_dir, f"{feature_file}.markers"))
        wav = wav[int(sent_begin * fs): int(sent_end * fs)]
        tgt_shape = (int(hp.n_mel_channels), hifi_mel_frames(len(wav), hp))
        if tgt_shape != tuple(mel_spec_pred.shape):
            with self._lock:
                self.n_skipped_shape += 1
            _logger.warning(f"{feature_dir} -- {feature_file} -- shape mismatch tgt {tgt_shape} pred {tuple(mel_spec_pred.shape)}, skipping")
            return
        if len(wav) < fs:
            with self._lock:
                self.n_skipped_short += 1
            _logger.warning(f"{feature_dir} -- {feature_file} -- Ignoring because audio is < 1s")
            return
        from scipy.io.wavfile import write
        wav = (wav * 32768.0).astype("int16")
        mel_spec_file = os.path.join(self.ft_data_set, speaker_name, f"{feature_file}.npy")
        out_wav_file = os.path.join(self.ft_data_set, speaker_name, f"{feature_file}.wav")
        try:
            # the crop is a view of a pinned staging buffer that the runner recycles: np.save copies it out
            np.save(mel_spec_file, np.ascontiguousarray(mel_spec_pred))
            write(out_wav_file, fs, wav)
            with self._lock:
                self.n_written += 1
        except Exception as e:          # noqa: BLE001 - the reference logs and removes the partial files (fine_tune.py:116-125)
            _logger.error(f"{feature_dir} -- {feature_file} -- {e}")
            if os.path.isfile(mel_spec_file):
                os.remove(mel_spec_file)
            if os.path.isfile(out_wav_file):
                os.remove(out_wav_file)

    def summary(self):
        return (f"Fine-tuning dataset summary: written={self.n_written}, skipped (shape mismatch)={self.n_skipped_shape}, "
                f"skipped (<1s)={self.n_skipped_short}")


def generate_fine_tuning_dataset(model, hparams, utterances, items, batch_size=64, rank=0, world=1, wav_loader=None,
                                 ft_data_set=None, **runner_kw):
    """``fine_tuning``'s main loop (``fine_tune.py:62-129``) on this rank's share of the utterances: the deterministic
    length-bucketed plan of ``bulk.plan_shards``, the pipelined ``bulk.BulkRunner`` and a ``FineTuneWriter`` as its sink.
    ``utterances``: the in-memory stream ``bulk`` takes (dict of per-utterance arrays, or a ``bulk.PointerTable``);
    ``items[i] = (feature_dir, feature_file)``.  Returns the writer (counters, ``summary()``)."""
    from . import bulk
    in_lens = [int(x.shape[0]) for x in utterances["symbols"]]
    out_lens = [int(x.shape[1]) for x in utterances["mel"]]
    batches, mine = bulk.plan_shards(in_lens, out_lens, batch_size, world)
    writer = FineTuneWriter(hparams, items, ft_data_set=ft_data_set, wav_loader=wav_loader)
    bulk.BulkRunner(model, sink=writer, **runner_kw).run(utterances, batches, mine[rank])
    _logger.info(writer.summary())
    return writer
