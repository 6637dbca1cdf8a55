"""TEST INFRASTRUCTURE ONLY - CPU restatement of the synthesis driver's tensor plumbing.

Follows ``/root/reference/src/daft_exprt/generate.py`` step by step, one utterance at a time, with plain loops:

  collate                 generate.py:104-162   sort by symbol count (descending), pad, permute names / speaker ids /
                                                external-prosody entries
  normalize_feature       generate.py:165-185   source -> target z-score mapping that keeps zeros
  external_tensors        generate.py:213-278   duration variance exaggeration, seconds + rounded frames, energy / pitch
                                                normalisation and alpha scaling
  generate_batch          generate.py:187-315   the three above + ``inference`` (the oracle port of model.py:1026-1114)
                                                + per-utterance crops; the vocoder / plot / file part (:317-330) is not
                                                restated
  generate                generate.py:333-381   chunk into batches; external embeddings are cut by position and are NOT
                                                permuted by the sort inside ``collate`` (that is what the reference does)

PINNED: ``tests/test_generate.py`` checks this file against the live reference's ``collate_tensors`` /
``generate_batch_mel_specs`` / ``generate_mel_specs`` (build container, through ``oracle/ref_harness.py``) and against the
frozen ``tests/golden/gen_*.npz`` made by ``tests/golden/make_generate_golden.py``.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module.
"""
import numpy as np
import torch

from oracle import daft_oracle as O


def collate(sentences, dur_factors, energy_factors, pitch_factors, pitch_transform, speaker_ids, file_names, hparams,
            external_prosody=None):
    """generate.py:104-162.  Returns the reference's 8-tuple."""
    items = []
    for sentence, df, ef, pf in zip(sentences, dur_factors, energy_factors, pitch_factors):
        ids = []
        for item in sentence:                               # a word is a list of phones, a boundary a single symbol
            if isinstance(item, list):
                for phone in item:
                    ids.append(hparams.symbols.index(phone))
            else:
                ids.append(hparams.symbols.index(item))
        n = len(ids)
        df = [1.0] * n if df is None else list(df)
        ef = [1.0] * n if ef is None else list(ef)
        if pf is None:
            pf = [0.0 if pitch_transform == "add" else 1.0] * n
        pf = list(pf)
        assert len(df) == n and len(ef) == n and len(pf) == n
        items.append((ids, df, ef, pf))
    lengths = torch.tensor([len(it[0]) for it in items], dtype=torch.long)
    input_lengths, order = torch.sort(lengths, dim=0, descending=True)      # the reference's sort (its tie order too)
    B, L = len(items), int(input_lengths[0])
    symbols = torch.zeros(B, L, dtype=torch.long)
    dur = torch.ones(B, L)
    energy = torch.ones(B, L)
    pitch = torch.zeros(B, L) if pitch_transform == "add" else torch.ones(B, L)
    names, spk = [], torch.zeros(B, dtype=torch.long)
    for i, src in enumerate(order.tolist()):
        ids, df, ef, pf = items[src]
        n = len(ids)
        symbols[i, :n] = torch.tensor(ids, dtype=torch.long)
        dur[i, :n] = torch.tensor(df, dtype=torch.float32)
        energy[i, :n] = torch.tensor(ef, dtype=torch.float32)
        pitch[i, :n] = torch.tensor(pf, dtype=torch.float32)
        names.append(file_names[src])
        spk[i] = int(speaker_ids[src])
    sorted_external = None
    if external_prosody is not None:
        assert len(external_prosody) == len(sentences)
        sorted_external = [external_prosody[src] for src in order.tolist()]
    return symbols, dur, energy, pitch, input_lengths, spk, names, sorted_external


def normalize_feature(values, zero_mask, target_stats, source_stats=None):
    """generate.py:165-185 in fp32 (torch promotes the Python-float statistics to the tensor's dtype): every non-zero
    element goes source z-score -> target scale -> target z-score; zeros stay zeros."""
    if source_stats is not None and source_stats["std"] == 0:
        raise ValueError("Source stats std cannot be 0.")
    if target_stats["std"] == 0:
        raise ValueError("Target speaker stats std cannot be 0.")
    v = values.clone()
    if source_stats is not None:
        v = (v - source_stats["mean"]) / source_stats["std"]
        v = v * target_stats["std"] + target_stats["mean"]
    v = (v - target_stats["mean"]) / target_stats["std"]
    return torch.where(zero_mask, torch.zeros_like(v), v)


def speaker_stats(hparams, speaker_id):
    """generate.py:244-255: stats of the speaker, falling back to 'spk 0'."""
    key = f"spk {speaker_id}"
    if key not in hparams.stats and "spk 0" in hparams.stats:
        key = "spk 0"
    if key not in hparams.stats:
        raise KeyError(f"Speaker stats missing for 'spk {speaker_id}' and fallback 'spk 0' not in hparams.stats "
                       f"(keys: {list(hparams.stats.keys())})")
    st = hparams.stats[key]
    if st["energy"]["std"] == 0 or st["pitch"]["std"] == 0:
        raise ValueError(f"Speaker stats not initialized for speaker ID {speaker_id}.")
    return st


def external_tensors(sorted_external, input_lengths, speaker_ids, hparams, max_len, source_stats=None,
                     alpha_dur=1.0, alpha_pitch=1.0, alpha_energy=1.0):
    """generate.py:213-278.  Returns the dict handed to ``inference`` (CPU tensors)."""
    B = len(sorted_external)
    dur = torch.zeros(B, max_len)
    dint = torch.zeros(B, max_len, dtype=torch.long)
    energy = torch.zeros(B, max_len)
    pitch = torch.zeros(B, max_len)
    hop_s = hparams.hop_length / hparams.sampling_rate
    for i, (entry, n) in enumerate(zip(sorted_external, input_lengths.tolist())):
        assert len(entry["symbols"]) == n
        frames = torch.tensor(list(entry["durations_frames"]), dtype=torch.float32)
        voiced = frames > 0
        if bool(voiced.any()) and alpha_dur != 1.0:
            mean = frames[voiced].mean()
            frames[voiced] = mean + alpha_dur * (frames[voiced] - mean)
            frames = torch.clamp(frames, min=0.0)
        dur[i, :n] = frames * hop_s
        dint[i, :n] = torch.round(frames).long()
        e = torch.tensor(list(entry["energy"]), dtype=torch.float32)
        p = torch.tensor(list(entry["pitch"]), dtype=torch.float32)
        e_zero, p_zero = e == 0.0, p == 0.0
        st = speaker_stats(hparams, int(speaker_ids[i]))
        e = normalize_feature(e, e_zero, dict(st["energy"]), None if source_stats is None else source_stats["energy"])
        p = normalize_feature(p, p_zero, dict(st["pitch"]), None if source_stats is None else source_stats["pitch"])
        if alpha_energy != 1.0:
            e[~e_zero] *= alpha_energy
        if alpha_pitch != 1.0:
            p[~p_zero] *= alpha_pitch
        energy[i, :n] = e
        pitch[i, :n] = p
    return {"duration_preds": dur, "durations_int": dint, "energy_preds": energy, "pitch_preds": pitch}


def generate_batch(state_dict, hparams, sentences, dur_factors, energy_factors, pitch_factors, pitch_transform,
                   speaker_ids, file_names, external_prosody=None, source_stats=None, alpha_dur=1.0, alpha_pitch=1.0,
                   alpha_energy=1.0, external_embeddings=None, external_accent_emb=None):
    """generate.py:187-315 without files: ``{name: [duration, duration_int, energy, pitch, mel, weight]}`` (numpy)."""
    names = [f"{n}_spk_{s}" for n, s in zip(file_names, speaker_ids)]                       # :195-197
    symbols, df, ef, pf, in_len, spk, names, sorted_ext = collate(
        sentences, dur_factors, energy_factors, pitch_factors, pitch_transform, speaker_ids, names, hparams,
        external_prosody)
    ext = None
    if sorted_ext is not None:
        ext = external_tensors(sorted_ext, in_len, spk, hparams, symbols.shape[1], source_stats, alpha_dur, alpha_pitch,
                               alpha_energy)
    enc, dec, weights = O.inference(state_dict, hparams, (symbols, df, ef, pf, in_len, spk), pitch_transform, ext,
                                    external_embeddings, external_accent_emb)
    d, di, e, p, in_len = [np.asarray(t) for t in enc]
    mel, out_len = np.asarray(dec[0]), np.asarray(dec[1])
    w = np.asarray(weights)
    preds = {}
    for i, name in enumerate(names):                                                         # :301-315
        n, t = int(in_len[i]), int(out_len[i])
        preds[name] = [d[i, :n], di[i, :n], e[i, :n], p[i, :n], mel[i, :, :t], w[i, :n, :t]]
    return preds


def generate(state_dict, hparams, sentences, file_names, speaker_ids, dur_factors=None, energy_factors=None,
             pitch_factors=None, batch_size=1, external_prosody=None, source_stats=None, alpha_dur=None, alpha_pitch=None,
             alpha_energy=None, external_embeddings=None, external_accent_emb=None):
    """generate.py:333-381: positional chunks of ``batch_size`` utterances, predictions merged into one dict."""
    n = len(sentences)
    dur_factors = [None] * n if dur_factors is None else dur_factors
    energy_factors = [None] * n if energy_factors is None else energy_factors
    pitch_factors = ["add", [None] * n] if pitch_factors is None else pitch_factors
    pitch_transform, pitch_factors = pitch_factors[0].lower(), pitch_factors[1]
    assert pitch_transform in ("add", "multiply")
    preds = {}
    for a in range(0, n, batch_size):
        b = min(a + batch_size, n)
        preds.update(generate_batch(
            state_dict, hparams, sentences[a:b], dur_factors[a:b], energy_factors[a:b], pitch_factors[a:b],
            pitch_transform, speaker_ids[a:b], list(file_names[a:b]),
            None if external_prosody is None else external_prosody[a:b], source_stats,
            alpha_dur, alpha_pitch, alpha_energy,
            None if external_embeddings is None else external_embeddings[a:b],
            None if external_accent_emb is None else external_accent_emb[a:b]))
    return preds
