"""Synthesis driver: the tensor plumbing on either side of ``DaftExprt.inference``.

Mirrors the reference's ``src/daft_exprt/generate.py`` for the part that touches tensors - same function names, argument
meaning, return values and error behaviour:

  collate_tensors               generate.py:104-162   sort by symbol count, pad, permute names / speaker ids / prosody
  _normalize_external_feature   generate.py:165-185   source -> target z-score mapping that keeps zeros
  generate_batch_mel_specs      generate.py:187-330   collate, external prosody tensors (:213-278), ``inference``, crops
  generate_mel_specs            generate.py:333-399   batches of ``batch_size`` utterances, optional RTF report

Text front end (phonemiser, ``generate.py:27-101``) and plotting (``:324-326``) are outside this row.  The HiFi-GAN
vocoder is row N4 (``vocoder.py``); any vocoder object handed in by the caller is used exactly
as the reference uses it (``:327-329``).

What is different from the reference is how the bytes move (the results are identical, tests/test_generate.py):

  * the reference looks every phone up with ``list.index`` and fills the padded tensors element by element
    (``:110-150``), then moves 10 pageable tensors with 10 ``.cuda()`` calls (``:204-210,273-278``).  Here one batch is
    assembled by numpy row assignments straight into ONE pinned staging block (symbols, factors, external prosody, lengths,
    speaker ids, embeddings), which crosses PCIe as a single asynchronous copy; the device tensors handed to ``inference``
    are views of its device mirror;
  * external-prosody normalisation (``:222-272``) runs once over the flat concatenation of the batch, one pass per
    (speaker statistics) group, with the reference's own fp32 expressions - not one Python iteration per utterance with
    six boolean-mask gathers each;
  * the reference reads the results back with 8 blocking ``.cpu()`` calls from pageable memory (``:290-297``).  Here the
    outputs go to pinned buffers on a copy stream behind an event; ``generate_mel_specs`` finishes (crops, ``.npz``)
    batch i while the GPU already runs batch i+1.  Crops are numpy views of the pinned buffers (no host copies), which the
    caching pinned allocator recycles once the predictions are dropped.

The host side needs the GPU library only when a batch is actually run: ``collate_tensors`` and the prosody tensors are
plain CPU code and are tested without a GPU.
"""
import logging
import os
import time

import numpy as np
import torch

_logger = logging.getLogger(__name__)

_ALIGN = 256


# ------------------------------------------------------------------------------------------------------
# staging block: every per-batch input in one allocation
# ------------------------------------------------------------------------------------------------------
class _Stage:
    """One byte block that holds all inputs of a batch; ``self.t[name]`` are typed torch views, ``self.a[name]`` the
    numpy views of the same memory.  Pinned when ``pinned`` (a CUDA device will read it)."""

    def __init__(self, B, L, emb_dim, acc_dim, pinned, with_prosody=True):
        i8, f4 = torch.int64, torch.float32
        plan = [("symbols", i8, (B, L)), ("input_lengths", i8, (B,)), ("speaker_ids", i8, (B,)),
                ("dur_factors", f4, (B, L)), ("energy_factors", f4, (B, L)), ("pitch_factors", f4, (B, L))]
        if with_prosody:
            plan += [("durations_int", i8, (B, L)), ("duration_preds", f4, (B, L)), ("energy_preds", f4, (B, L)),
                     ("pitch_preds", f4, (B, L))]
        if emb_dim:
            plan.append(("external_embeddings", f4, (B, emb_dim)))
        if acc_dim:
            plan.append(("external_accent_emb", f4, (B, acc_dim)))
        self.layout, off = {}, 0
        for name, dt, shape in plan:
            nbytes = int(np.prod(shape)) * (8 if dt == i8 else 4)
            self.layout[name] = (off, nbytes, dt, shape)
            off += (nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
        self.nbytes = max(off, _ALIGN)
        self.block = torch.empty(self.nbytes, dtype=torch.uint8, pin_memory=bool(pinned))
        self.t = self.views(self.block)
        self.a = {k: v.numpy() for k, v in self.t.items()}

    def views(self, block):
        """Typed views of ``block`` (this block or its device mirror)."""
        return {name: block[off:off + nbytes].view(dt).view(shape) for name, (off, nbytes, dt, shape) in self.layout.items()}


_SYMBOL_TABLES = {}


def _symbol_table(hparams):
    """symbol -> first index in ``hparams.symbols`` (what ``list.index`` returns, generate.py:115-117)."""
    syms = hparams.symbols
    key = id(syms)
    hit = _SYMBOL_TABLES.get(key)
    if hit is None or hit[0] is not syms or hit[1] != len(syms):
        table = {}
        for i, s in enumerate(syms):
            table.setdefault(s, i)
        if len(_SYMBOL_TABLES) > 16:
            _SYMBOL_TABLES.clear()
        hit = (syms, len(syms), table)
        _SYMBOL_TABLES[key] = hit
    return hit[2]


def _sentence_ids(sentence, table):
    ids = []
    try:
        for item in sentence:
            if isinstance(item, list):              # a word: list of phones; anything else: one boundary symbol
                ids.extend([table[p] for p in item])
            else:
                ids.append(table[item])
    except KeyError as e:                          # list.index raises ValueError (generate.py:115,117)
        raise ValueError(f"{e.args[0]!r} is not in list") from None
    return ids


def _collate_into(stage_factory, batch_sentences, batch_dur_factors, batch_energy_factors, batch_pitch_factors,
                  pitch_transform, batch_speaker_ids, batch_file_names, hparams, external_prosody):
    table = _symbol_table(hparams)
    ids = [_sentence_ids(s, table) for s, _, _, _ in
           zip(batch_sentences, batch_dur_factors, batch_energy_factors, batch_pitch_factors)]
    n_items = len(ids)
    for k, (df, ef, pf) in enumerate(zip(batch_dur_factors, batch_energy_factors, batch_pitch_factors)):
        n = len(ids[k])
        assert df is None or len(df) == n
        assert ef is None or len(ef) == n
        assert pf is None or len(pf) == n
    # the reference's own sort call (generate.py:135-136), so ties come out in the same order
    input_lengths, order = torch.sort(torch.LongTensor([len(x) for x in ids]), dim=0, descending=True)
    order = order.tolist()
    B, L = n_items, int(input_lengths[0])
    stage = stage_factory(B, L)
    a = stage.a
    a["symbols"].fill(0)
    a["dur_factors"].fill(1.0)
    a["energy_factors"].fill(1.0)
    a["pitch_factors"].fill(0.0 if pitch_transform == "add" else 1.0)
    a["input_lengths"][:] = input_lengths.numpy()
    for i, src in enumerate(order):
        n = len(ids[src])
        a["symbols"][i, :n] = ids[src]
        if batch_dur_factors[src] is not None:
            a["dur_factors"][i, :n] = batch_dur_factors[src]
        if batch_energy_factors[src] is not None:
            a["energy_factors"][i, :n] = batch_energy_factors[src]
        if batch_pitch_factors[src] is not None:
            a["pitch_factors"][i, :n] = batch_pitch_factors[src]
        a["speaker_ids"][i] = batch_speaker_ids[src]
    file_names = [batch_file_names[src] for src in order]
    sorted_external = None
    if external_prosody is not None:
        assert len(external_prosody) == len(batch_sentences)
        sorted_external = [external_prosody[src] for src in order]
    return stage, file_names, sorted_external


def collate_tensors(batch_sentences, batch_dur_factors, batch_energy_factors, batch_pitch_factors, pitch_transform,
                    batch_speaker_ids, batch_file_names, hparams, external_prosody=None):
    """``collate_tensors`` (reference generate.py:104-162); same 8-tuple:
    ``(symbols, dur_factors, energy_factors, pitch_factors, input_lengths, speaker_ids, file_names, sorted_external)``."""
    stage, file_names, sorted_external = _collate_into(
        lambda B, L: _Stage(B, L, 0, 0, pinned=False, with_prosody=False), batch_sentences, batch_dur_factors,
        batch_energy_factors, batch_pitch_factors, pitch_transform, batch_speaker_ids, batch_file_names, hparams,
        external_prosody)
    t = stage.t
    return (t["symbols"], t["dur_factors"], t["energy_factors"], t["pitch_factors"], t["input_lengths"],
            t["speaker_ids"], file_names, sorted_external)


# ------------------------------------------------------------------------------------------------------
# external prosody (reference generate.py:165-185, 213-278)
# ------------------------------------------------------------------------------------------------------
def _zscore_map(values, sel, target_stats, source_stats):
    """The reference's fp32 expressions (generate.py:172-183) on the elements ``sel`` of the 1-D tensor ``values``."""
    if source_stats is not None:
        src_mean = source_stats['mean']
        src_std = source_stats['std']
        if src_std == 0:
            raise ValueError('Source stats std cannot be 0.')
        tmp = values[sel]
        tmp = (tmp - src_mean) / src_std
        tmp = tmp * target_stats['std'] + target_stats['mean']
        values[sel] = tmp
    tgt_std = target_stats['std']
    if tgt_std == 0:
        raise ValueError('Target speaker stats std cannot be 0.')
    tmp = values[sel]
    tmp = (tmp - target_stats['mean']) / tgt_std
    values[sel] = tmp


def _normalize_external_feature(values, zero_mask, target_stats, source_stats=None):
    """``_normalize_external_feature`` (reference generate.py:165-185): source -> target z-score mapping, zeros kept."""
    values = values.clone()
    _zscore_map(values, ~zero_mask, target_stats, source_stats)
    values[zero_mask] = 0.
    return values


def _speaker_stats(hparams, speaker_id):
    """Statistics of one speaker with the reference's fallback and errors (generate.py:244-255)."""
    spk_key = f'spk {speaker_id}'
    if spk_key not in hparams.stats and 'spk 0' in hparams.stats:
        spk_key = 'spk 0'
    if spk_key not in hparams.stats:
        raise KeyError(f"Speaker stats missing for 'spk {speaker_id}' and fallback 'spk 0' not in hparams.stats "
                       f"(keys: {list(hparams.stats.keys())})")
    st = hparams.stats[spk_key]
    if st['energy']['std'] == 0 or st['pitch']['std'] == 0:
        raise ValueError(f'Speaker stats not initialized for speaker ID {speaker_id}.')
    return spk_key, st


def _flat_f32(entries, key, lens):
    parts = [np.asarray(e[key], dtype=np.float32).reshape(-1) for e in entries]
    for p, n in zip(parts, lens):
        if p.shape[0] != n:                        # the reference fails on the padded-row assignment (generate.py:238,271)
            raise RuntimeError(f"external prosody '{key}' has {p.shape[0]} values for {n} symbols")
    return torch.from_numpy(np.concatenate(parts) if parts else np.zeros(0, np.float32))


def _external_prosody_into(a, sorted_external, input_lengths, speaker_ids, file_names, hparams, source_stats=None,
                           alpha_dur=1.0, alpha_pitch=1.0, alpha_energy=1.0):
    """Fill the planes ``duration_preds / durations_int / energy_preds / pitch_preds`` of ``a`` (numpy views, (B, L_max))
    from the sorted prosody entries: reference generate.py:213-272 over the flat concatenation of the batch."""
    lens = [int(n) for n in input_lengths]
    B = len(lens)
    src_energy = source_stats['energy'] if source_stats is not None else None
    src_pitch = source_stats['pitch'] if source_stats is not None else None
    groups, group_of = {}, []
    for idx, (entry, n) in enumerate(zip(sorted_external, lens)):          # errors in the reference's order
        assert len(entry['symbols']) == n, \
            _logger.error(f'External prosody length mismatch for sample {file_names[idx]}')
        key, st = _speaker_stats(hparams, int(speaker_ids[idx]))
        if idx == 0:
            for src in (src_energy, src_pitch):
                if src is not None and src['std'] == 0:
                    raise ValueError('Source stats std cannot be 0.')
        group_of.append(groups.setdefault(key, (len(groups), st))[0])
    frames = _flat_f32(sorted_external, 'durations_frames', lens)
    energy = _flat_f32(sorted_external, 'energy', lens)
    pitch = _flat_f32(sorted_external, 'pitch', lens)
    starts = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)

    # variance exaggeration of the durations: per-utterance mean of the voiced frames (generate.py:228-233)
    if alpha_dur != 1.0:
        for i in range(B):
            fr = frames[starts[i]:starts[i + 1]]
            dur_mask = fr > 0
            if dur_mask.any():
                dur_mean = fr[dur_mask].mean()
                fr[dur_mask] = dur_mean + alpha_dur * (fr[dur_mask] - dur_mean)
                fr.clamp_(min=0.0)
    hop_in_seconds = hparams.hop_length / hparams.sampling_rate
    duration = frames * hop_in_seconds
    duration_int = torch.round(frames).long()

    energy_zero = energy == 0.
    pitch_zero = pitch == 0.
    if len(groups) == 1:
        st = next(iter(groups.values()))[1]
        _zscore_map(energy, ~energy_zero, st['energy'], src_energy)
        _zscore_map(pitch, ~pitch_zero, st['pitch'], src_pitch)
    else:
        elem_group = torch.from_numpy(np.repeat(np.asarray(group_of, dtype=np.int64), lens))
        for gid, st in groups.values():
            in_group = elem_group == gid
            _zscore_map(energy, in_group & ~energy_zero, st['energy'], src_energy)
            _zscore_map(pitch, in_group & ~pitch_zero, st['pitch'], src_pitch)
    energy[energy_zero] = 0.
    pitch[pitch_zero] = 0.
    if alpha_energy != 1.0:
        energy[~energy_zero] *= alpha_energy
    if alpha_pitch != 1.0:
        pitch[~pitch_zero] *= alpha_pitch

    rows = np.repeat(np.arange(B), lens)
    cols = np.arange(int(starts[-1])) - np.repeat(starts[:-1], lens)
    for name, flat in (("duration_preds", duration), ("durations_int", duration_int), ("energy_preds", energy),
                       ("pitch_preds", pitch)):
        a[name].fill(0)
        a[name][rows, cols] = flat.numpy()


def external_prosody_tensors(sorted_external_prosody, input_lengths, speaker_ids, file_names, hparams, max_len,
                             source_stats=None, alpha_dur=1.0, alpha_pitch=1.0, alpha_energy=1.0):
    """The ``external_tensors`` dict of the reference (generate.py:213-278) as CPU tensors."""
    B = len(sorted_external_prosody)
    a = {"duration_preds": np.zeros((B, max_len), np.float32), "durations_int": np.zeros((B, max_len), np.int64),
         "energy_preds": np.zeros((B, max_len), np.float32), "pitch_preds": np.zeros((B, max_len), np.float32)}
    _external_prosody_into(a, sorted_external_prosody, input_lengths, speaker_ids, file_names, hparams, source_stats,
                           alpha_dur, alpha_pitch, alpha_energy)
    return {k: torch.from_numpy(v) for k, v in a.items()}


# ------------------------------------------------------------------------------------------------------
# one batch: launch (host staging -> one H2D -> inference -> async D2H) and finish (event wait, crops, files)
# ------------------------------------------------------------------------------------------------------
class _PendingBatch:
    """A batch whose results are on their way to pinned host memory."""

    def __init__(self, file_names, input_lengths, host, done, keep, h2d_bytes, d2h_bytes):
        self.file_names, self.input_lengths = file_names, input_lengths
        self.host, self.done, self._keep = host, done, keep
        self.h2d_bytes, self.d2h_bytes = h2d_bytes, d2h_bytes

    def finish(self):
        """Wait for the copies and cut the per-utterance views (reference generate.py:299-315).
        ``{file_name: [duration_pred, duration_int, energy_pred, pitch_pred, mel_spec_pred, weight]}``"""
        self.done.synchronize()
        self._keep = None                                    # device tensors may go back to the allocator now
        h = self.host
        prosody = h["prosody"].numpy()                       # (3, B, L_max): duration, energy, pitch
        durations_int = h["durations_int"].numpy()
        mel = h["mel"].numpy()                               # (B, n_mel_channels, T_max)
        weights = h["weights"].numpy()                       # (B, L_max, T_max)
        in_len = self.input_lengths
        out_len = h["output_lengths"].numpy()
        predictions = {}
        for i, name in enumerate(self.file_names):
            n, t = int(in_len[i]), int(out_len[i])
            predictions[f'{name}'] = [prosody[0, i, :n], durations_int[i, :n], prosody[1, i, :n], prosody[2, i, :n],
                                      mel[i, :, :t], weights[i, :n, :t]]
        return predictions


_COPY_STREAMS = {}


def _copy_stream(device):
    key = (device.type, device.index)
    if key not in _COPY_STREAMS:
        _COPY_STREAMS[key] = torch.cuda.Stream(device)
    return _COPY_STREAMS[key]


def _embedding_rows(x, B, what):
    """Rows of an embedding argument as (tensor-or-None, width).  The reference hands them to ``inference`` untouched
    (generate.py:280-284): cut by position, NOT permuted by the length sort."""
    if x is None:
        return None, 0
    if not torch.is_tensor(x):
        x = torch.as_tensor(np.asarray(x))
    if x.dim() != 2 or x.shape[0] != B:
        raise ValueError(f"{what} must be (batch, dim) = ({B}, dim), got {tuple(x.shape)}")
    return x, int(x.shape[1])


def launch_batch_mel_specs(model, batch_sentences, batch_dur_factors, batch_energy_factors, batch_pitch_factors,
                           pitch_transform, batch_speaker_ids, batch_file_names, hparams, batch_external_prosody=None,
                           source_stats=None, alpha_dur=1.0, alpha_pitch=1.0, alpha_energy=1.0,
                           external_embeddings=None, external_accent_emb=None):
    """Everything of ``generate_batch_mel_specs`` up to and including the device-to-host copies, without waiting for
    them (reference generate.py:195-297).  Returns a ``_PendingBatch``; ``.finish()`` gives the predictions."""
    device = model.device
    if device.type != "cuda":
        raise RuntimeError("daft_exprt_b200.generate runs the B200 kernels: move the model to a CUDA device first")
    for idx, file_name in enumerate(batch_file_names):                      # generate.py:195-197 (in place, like there)
        batch_file_names[idx] = file_name + f'_spk_{batch_speaker_ids[idx]}'
        _logger.info(f'Generating "{batch_sentences[idx]}" as "{batch_file_names[idx]}"')
    B = len(batch_sentences)
    emb, emb_dim = _embedding_rows(external_embeddings, B, "external_embeddings")
    acc, acc_dim = _embedding_rows(external_accent_emb, B, "external_accent_emb")
    stage_emb = emb_dim if emb is not None and emb.device.type == "cpu" else 0
    stage_acc = acc_dim if acc is not None and acc.device.type == "cpu" else 0
    has_prosody = batch_external_prosody is not None
    stage, file_names, sorted_external = _collate_into(
        lambda b, l: _Stage(b, l, stage_emb, stage_acc, pinned=True, with_prosody=has_prosody), batch_sentences,
        batch_dur_factors, batch_energy_factors, batch_pitch_factors, pitch_transform, batch_speaker_ids,
        batch_file_names, hparams, batch_external_prosody)
    a = stage.a
    input_lengths = a["input_lengths"].copy()
    if has_prosody:
        _external_prosody_into(a, sorted_external, input_lengths, a["speaker_ids"], file_names, hparams, source_stats,
                               alpha_dur, alpha_pitch, alpha_energy)
    if stage_emb:
        stage.t["external_embeddings"].copy_(emb)
    if stage_acc:
        stage.t["external_accent_emb"].copy_(acc)

    # one copy for the whole batch; the device tensors are views of the mirror block
    mirror = torch.empty(stage.nbytes, dtype=torch.uint8, device=device)
    mirror.copy_(stage.block, non_blocking=True)
    d = stage.views(mirror)
    inputs = (d["symbols"], d["dur_factors"], d["energy_factors"], d["pitch_factors"], d["input_lengths"], d["speaker_ids"])
    external_tensors = None
    if has_prosody:
        external_tensors = {k: d[k] for k in ("duration_preds", "durations_int", "energy_preds", "pitch_preds")}
    encoder_preds, decoder_preds, alignments = model.inference(
        inputs, pitch_transform, hparams, external_tensors,
        external_embeddings=d["external_embeddings"] if stage_emb else emb,
        external_accent_emb=d["external_accent_emb"] if stage_acc else acc)
    duration_preds, durations_int, energy_preds, pitch_preds, _ = encoder_preds
    mel_spec_preds, output_lengths = decoder_preds

    # results -> pinned memory on the copy stream, behind the compute stream
    compute = torch.cuda.current_stream(device)
    prosody = torch.stack([duration_preds, energy_preds, pitch_preds])       # (3, B, L_max), one small kernel
    ready = compute.record_event()
    side = _copy_stream(device)
    host = {"prosody": torch.empty(prosody.shape, dtype=torch.float32, pin_memory=True),
            "durations_int": torch.empty(durations_int.shape, dtype=torch.int64, pin_memory=True),
            "output_lengths": torch.empty(output_lengths.shape, dtype=torch.int64, pin_memory=True),
            "mel": torch.empty(mel_spec_preds.shape, dtype=torch.float32, pin_memory=True),
            "weights": torch.empty(alignments.shape, dtype=torch.float32, pin_memory=True)}
    dev = {"prosody": prosody, "durations_int": durations_int, "output_lengths": output_lengths, "mel": mel_spec_preds,
           "weights": alignments}
    with torch.cuda.stream(side):
        side.wait_event(ready)
        for k in ("output_lengths", "prosody", "durations_int", "mel", "weights"):
            host[k].copy_(dev[k], non_blocking=True)
        done = side.record_event()
    d2h = sum(t.numel() * t.element_size() for t in host.values())
    return _PendingBatch(file_names, input_lengths, host, done, (dev, mirror, stage), stage.nbytes, d2h)


def _write_batch(predictions, output_dir, hparams, vocoder):
    """Files of one batch (reference generate.py:311-312, 317-329): mel ``.npz`` always, wave when there is a vocoder."""
    for file_name, preds in predictions.items():
        np.savez(os.path.join(output_dir, f'{file_name}.npz'), mel_spec=preds[4])
    if vocoder is None:
        raise ValueError(
            "HiFi-GAN vocoder required for mel-to-wave. "
            "Provide --vocoder_checkpoint in synthesize.py or load a vocoder before calling generate."
        )
    from scipy.io import wavfile
    names = list(predictions)
    audios = None
    if hasattr(vocoder, "infer_batch") and len(names) > 1:
        # ours (vocoder.HiFiGanVocoder): the whole batch in ONE generator pass over the packed frame stream, every
        # utterance vocoded as if alone (bit-identical to per-utterance infer calls, tests/test_gpu_vocoder.py)
        mels = [np.asarray(predictions[n][4], dtype=np.float32) for n in names]
        lengths = [m.shape[1] for m in mels]
        batch = np.zeros((len(mels), mels[0].shape[0], max(lengths)), dtype=np.float32)
        for b, m in enumerate(mels):
            batch[b, :, : m.shape[1]] = m
        audios = dict(zip(names, vocoder.infer_batch(batch, lengths)))
    for file_name, (_, _, _, _, mel_spec, weight) in predictions.items():
        audio = audios[file_name] if audios is not None else vocoder.infer(mel_spec)
        audio_int16 = (audio * 32767.5).clip(min=-32768, max=32767).astype(np.int16)
        wavfile.write(os.path.join(output_dir, f'{file_name}.wav'), hparams.sampling_rate, audio_int16)


def generate_batch_mel_specs(model, batch_sentences, batch_dur_factors,
                             batch_energy_factors, batch_pitch_factors, pitch_transform,
                             batch_speaker_ids, batch_file_names, output_dir, hparams,
                             n_jobs, batch_external_prosody=None,
                             vocoder=None, source_stats=None,
                             alpha_dur=1.0, alpha_pitch=1.0, alpha_energy=1.0,
                             external_embeddings=None, external_accent_emb=None):
    """``generate_batch_mel_specs`` (reference generate.py:187-330): same arguments, same ``predictions`` dict, same
    files except the ``.png`` plots.  Like the reference it raises ``ValueError`` after the ``.npz`` files are written when
    no vocoder is given; ``launch_batch_mel_specs(...).finish()`` is the tensor-only call."""
    pending = launch_batch_mel_specs(
        model, batch_sentences, batch_dur_factors, batch_energy_factors, batch_pitch_factors, pitch_transform,
        batch_speaker_ids, batch_file_names, hparams, batch_external_prosody, source_stats, alpha_dur, alpha_pitch,
        alpha_energy, external_embeddings, external_accent_emb)
    predictions = pending.finish()
    _write_batch(predictions, output_dir, hparams, vocoder)
    return predictions


def _chunks(seq, size):
    return [seq[pos: pos + size] for pos in range(0, len(seq), size)]


def generate_mel_specs(model, sentences, file_names, speaker_ids, output_dir, hparams,
                       dur_factors=None, energy_factors=None, pitch_factors=None, batch_size=1,
                       n_jobs=1, get_time_perf=False, external_prosody=None,
                       vocoder=None, source_stats=None,
                       alpha_dur=None, alpha_pitch=None, alpha_energy=None,
                       external_embeddings=None, external_accent_emb=None):
    """``generate_mel_specs`` (reference generate.py:333-399): same arguments and result.

    Batches are pipelined two deep: batch i is cropped and written while the GPU runs batch i+1.  ``output_dir=None``
    skips every file (and with it the vocoder requirement): predictions only."""
    dur_factors = [None for _ in range(len(sentences))] if dur_factors is None else dur_factors
    energy_factors = [None for _ in range(len(sentences))] if energy_factors is None else energy_factors
    pitch_factors = ['add', [None for _ in range(len(sentences))]] if pitch_factors is None else pitch_factors
    pitch_transform = pitch_factors[0].lower()
    pitch_factors = pitch_factors[1]
    assert pitch_transform in ['add', 'multiply']
    assert len(file_names) == len(sentences)
    assert len(speaker_ids) == len(sentences)
    assert len(dur_factors) == len(sentences)
    assert len(energy_factors) == len(sentences)
    assert len(pitch_factors) == len(sentences)
    if external_prosody is not None:
        assert len(external_prosody) == len(sentences)

    model.eval()
    if output_dir is not None:
        os.makedirs(output_dir, exist_ok=True)
    n_chunks = len(_chunks(sentences, batch_size))
    chunk = lambda x: _chunks(x, batch_size) if x is not None else [None] * n_chunks     # noqa: E731
    sentence_chunks, dur_chunks, energy_chunks, pitch_chunks = chunk(sentences), chunk(dur_factors), chunk(energy_factors), chunk(pitch_factors)
    speaker_chunks, file_chunks, external_chunks = chunk(speaker_ids), chunk(list(file_names)), chunk(external_prosody)
    emb_chunks, accent_chunks = chunk(external_embeddings), chunk(external_accent_emb)

    predictions, time_per_batch = {}, []
    begin = time.time()

    def finish(pending):
        batch_predictions = pending.finish()
        if output_dir is not None:
            _write_batch(batch_predictions, output_dir, hparams, vocoder)
        predictions.update(batch_predictions)

    in_flight = None
    with torch.no_grad():
        for idx in range(n_chunks):
            pending = launch_batch_mel_specs(
                model, sentence_chunks[idx], dur_chunks[idx], energy_chunks[idx], pitch_chunks[idx], pitch_transform,
                speaker_chunks[idx], file_chunks[idx], hparams, external_chunks[idx], source_stats,
                alpha_dur, alpha_pitch, alpha_energy, emb_chunks[idx], accent_chunks[idx])
            if in_flight is not None:
                finish(in_flight)
            in_flight = pending
        if in_flight is not None:
            finish(in_flight)
    time_per_batch = [time.time() - begin]

    # display overall time performance (reference generate.py:383-397)
    if get_time_perf:
        durations = []
        for prediction in predictions.values():
            _, _, _, _, mel_spec, _ = prediction
            nb_frames = mel_spec.shape[1]
            nb_wav_samples = (nb_frames - 1) * hparams.hop_length + hparams.filter_length
            if hparams.centered:
                nb_wav_samples -= 2 * int(hparams.filter_length / 2)
            durations.append(nb_wav_samples / hparams.sampling_rate)
        _logger.info(f'{len(predictions)} sentences ({sum(durations):.2f}s) generated in {sum(time_per_batch):.2f}s')
        _logger.info(f'DaftExprt RTF: {sum(durations)/sum(time_per_batch):.2f}')

    return predictions
