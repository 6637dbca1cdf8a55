"""Seeded synthetic batches in the reference's tensor contract.

``forward_batch`` builds the 12-tuple ``DaftExprt.forward`` takes (reference ``model.py:893-896``):
rows sorted by symbol count descending and right-zero-padded, exactly as the reference's collate
does (``data_loader.py:217-263``), with ``sum(durations_int[b]) == output_lengths[b]``
(``data_loader.py:165-168``).  ``inference_batch`` builds the 6-tuple + external dicts of
``DaftExprt.inference`` (``model.py:1035``, ``generate.py:273-279``).  All CPU tensors.
"""
import torch


def _lengths(gen, batch, lmin, lmax):
    if lmin == lmax:
        lens = torch.full((batch,), lmin, dtype=torch.long)
    else:
        lens = torch.randint(lmin, lmax + 1, (batch,), generator=gen, dtype=torch.long)
    return torch.sort(lens, descending=True)[0]


def forward_batch(hp, batch, lmin, lmax, seed, dmin=1, dmax=9, const_dur=None):
    """12-tuple for ``forward``.  cfg1: batch=1,lmin=lmax=60,const_dur=5,seed=0.
    cfg2: batch=64,lmin=20,lmax=150,seed=42.  cfg4: lmin=lmax=400,const_dur=5,seed=9."""
    g = torch.Generator().manual_seed(int(seed))
    in_len = _lengths(g, batch, lmin, lmax)
    L = int(in_len.max())
    valid = torch.arange(L).unsqueeze(0) < in_len.unsqueeze(1)
    symbols = torch.randint(1, hp.n_symbols, (batch, L), generator=g, dtype=torch.long) * valid
    if const_dur is not None:
        dint = torch.full((batch, L), int(const_dur), dtype=torch.long) * valid
    else:
        dint = torch.randint(dmin, dmax + 1, (batch, L), generator=g, dtype=torch.long) * valid
    dfloat = dint.float() * (hp.hop_length / hp.sampling_rate)
    s_energy = torch.randn(batch, L, generator=g) * valid
    s_pitch = torch.randn(batch, L, generator=g) * valid
    out_len = dint.sum(dim=1)
    T = int(out_len.max())
    fvalid = torch.arange(T).unsqueeze(0) < out_len.unsqueeze(1)
    f_energy = torch.randn(batch, T, generator=g).abs() * fvalid
    f_pitch = torch.randn(batch, T, generator=g).abs() * fvalid
    mel = torch.randn(batch, hp.n_mel_channels, T, generator=g) * fvalid.unsqueeze(1)
    spk_ids = torch.randint(0, hp.n_speakers, (batch,), generator=g, dtype=torch.long)
    spk_embs = torch.randn(batch, getattr(hp, "external_emb_dim", 192), generator=g)
    return (symbols, dfloat, dint, s_energy, s_pitch, in_len, f_energy, f_pitch, mel, out_len,
            spk_ids, spk_embs)


def inference_batch(hp, batch, lmin, lmax, seed, pitch_transform="add"):
    """(inputs6, external_prosody, external_embeddings, external_accent_emb) for ``inference``.
    Durations are 1..9 hops (+ jitter) per valid symbol so every symbol gets ≥ 1 frame."""
    g = torch.Generator().manual_seed(int(seed))
    in_len = _lengths(g, batch, lmin, lmax)
    L = int(in_len.max())
    valid = torch.arange(L).unsqueeze(0) < in_len.unsqueeze(1)
    symbols = torch.randint(1, hp.n_symbols, (batch, L), generator=g, dtype=torch.long) * valid
    hops = torch.randint(1, 10, (batch, L), generator=g).float() + torch.rand(batch, L, generator=g) * 0.5
    dur = hops * (hp.hop_length / hp.sampling_rate) * valid
    energy = torch.randn(batch, L, generator=g) * valid
    pitch = torch.randn(batch, L, generator=g) * valid
    dur_f = torch.ones(batch, L)
    en_f = torch.ones(batch, L)
    pi_f = torch.zeros(batch, L) if pitch_transform == "add" else torch.ones(batch, L)
    spk_ids = torch.randint(0, hp.n_speakers, (batch,), generator=g, dtype=torch.long)
    ext_emb = torch.randn(batch, getattr(hp, "external_emb_dim", 192), generator=g)
    acc_emb = torch.randn(batch, hp.phoneme_encoder["hidden_embed_dim"], generator=g)
    inputs = (symbols, dur_f, en_f, pi_f, in_len, spk_ids)
    prosody = {"duration_preds": dur, "durations_int": torch.zeros(batch, L, dtype=torch.long),
               "energy_preds": energy, "pitch_preds": pitch}
    return inputs, prosody, ext_emb, acc_emb


CONFIGS = {
    # name: kwargs of forward_batch
    "cfg1": dict(batch=1, lmin=60, lmax=60, seed=0, const_dur=5),
    "cfg2": dict(batch=64, lmin=20, lmax=150, seed=42),
    "cfg4_b1": dict(batch=1, lmin=400, lmax=400, seed=9, const_dur=5),
    "cfg4_b8": dict(batch=8, lmin=400, lmax=400, seed=9, const_dur=5),
}


def config_batch(hp, name):
    return forward_batch(hp, **CONFIGS[name])


# ----------------------------------------------------------------------------------------------
# utterance streams (BASELINE.json config 3: bulk generation over many utterances)
# ----------------------------------------------------------------------------------------------
def _utt_lengths(hp, index, seed, lmin, lmax, dmin, dmax):
    g = torch.Generator().manual_seed(int(seed) + int(index))
    L = int(torch.randint(lmin, lmax + 1, (1,), generator=g))
    dint = torch.randint(dmin, dmax + 1, (L,), generator=g, dtype=torch.long)
    return g, L, dint


def stream_lengths(hp, n, seed=43, lmin=20, lmax=150, dmin=1, dmax=9):
    """(in_lens, out_lens) lists of utterances 0..n-1 of the stream, without materialising them."""
    in_lens, out_lens = [], []
    for i in range(n):
        _, L, dint = _utt_lengths(hp, i, seed, lmin, lmax, dmin, dmax)
        in_lens.append(L)
        out_lens.append(int(dint.sum()))
    return in_lens, out_lens


def utterance(hp, index, seed=43, lmin=20, lmax=150, dmin=1, dmax=9):
    """One utterance of the stream in the reference's per-item layout (``data_loader.py:123-178``)."""
    g, L, dint = _utt_lengths(hp, index, seed, lmin, lmax, dmin, dmax)
    T = int(dint.sum())
    return {
        "symbols": torch.randint(1, hp.n_symbols, (L,), generator=g, dtype=torch.long),
        "durations_int": dint,
        "durations_float": dint.float() * (hp.hop_length / hp.sampling_rate),
        "symbols_energy": torch.randn(L, generator=g),
        "symbols_pitch": torch.randn(L, generator=g),
        "frames_energy": torch.randn(T, generator=g).abs(),
        "frames_pitch": torch.randn(T, generator=g).abs(),
        "mel": torch.randn(hp.n_mel_channels, T, generator=g),
        "speaker_id": torch.randint(0, hp.n_speakers, (), generator=g, dtype=torch.long),
        "spk_emb": torch.randn(getattr(hp, "external_emb_dim", 192), generator=g),
    }


class UtteranceStream:
    """Lazy dict-of-lists view over a synthetic stream: ``stream[field][i]`` builds utterance i on demand
    (a small LRU keeps the ones of the batch being collated)."""

    FIELDS = ("symbols", "durations_int", "durations_float", "symbols_energy", "symbols_pitch", "frames_energy",
              "frames_pitch", "mel", "speaker_id", "spk_emb")

    class _Field:
        def __init__(self, owner, name):
            self.owner, self.name = owner, name

        def __getitem__(self, i):
            return self.owner._get(i)[self.name]

    def __init__(self, hp, n, seed=43, **kw):
        self.hp, self.n, self.seed, self.kw = hp, n, seed, kw
        self._cache = {}

    def _get(self, i):
        u = self._cache.get(i)
        if u is None:
            if len(self._cache) > 256:
                self._cache.clear()
            u = utterance(self.hp, i, self.seed, **self.kw)
            self._cache[i] = u
        return u

    def __getitem__(self, name):
        return UtteranceStream._Field(self, name)

    def lengths(self):
        return stream_lengths(self.hp, self.n, self.seed, **self.kw)


# ----------------------------------------------------------------------------------------------
# synthesis requests (the driver's inputs, reference generate.py:333-345)
# ----------------------------------------------------------------------------------------------
def synthesis_request(hp, n, lmin, lmax, seed, pitch_transform="add", with_factors=True, zero_frac=0.15):
    """``n`` utterances in the form ``generate_mel_specs`` takes them: phonemised sentences (words = lists of phones,
    boundaries = single symbols, EOS last; ``generate.py:60-77``), optional per-symbol factors, speaker ids, file names, the
    external prosody entries (``generate.py:222-243``: symbols, durations in frames, raw energy / log-pitch with zeros for
    silent / unvoiced symbols), ECAPA embeddings and accent embeddings.  Returned as keyword arguments of
    ``generate_mel_specs`` (minus model / output_dir / hparams)."""
    g = torch.Generator().manual_seed(int(seed))
    phones = [s for s in hp.symbols[7:]]
    sentences, prosody, dur_f, en_f, pi_f = [], [], [], [], []
    for i in range(n):
        L = int(torch.randint(lmin, lmax + 1, (1,), generator=g))
        sent, flat = [], []
        while len(flat) < L - 1:
            w = min(int(torch.randint(1, 6, (1,), generator=g)), L - 1 - len(flat))
            word = [phones[int(k)] for k in torch.randint(0, len(phones), (w,), generator=g)]
            sent.append(word)
            flat += word
            if len(flat) < L - 1:
                bound = " ,.!?"[int(torch.randint(0, 5, (1,), generator=g)) if int(torch.randint(0, 4, (1,), generator=g)) == 0 else 0]
                sent.append(bound)
                flat.append(bound)
        sent.append("~")
        flat.append("~")
        L = len(flat)
        silent = torch.tensor([s in " ,.!?~" for s in flat])
        frames = torch.randint(1, 10, (L,), generator=g).float() + torch.rand(L, generator=g) * 0.4
        drop = torch.rand(L, generator=g) < zero_frac
        frames = torch.where(silent & drop, torch.zeros(L), frames)
        energy = (torch.rand(L, generator=g) * 3.0 + 0.2) * (frames > 0)
        pitch = (5.0 + 0.3 * torch.randn(L, generator=g)) * (frames > 0) * (~silent) * (torch.rand(L, generator=g) > zero_frac)
        sentences.append(sent)
        prosody.append({"symbols": flat, "durations_frames": frames.tolist(), "energy": energy.tolist(),
                        "pitch": pitch.tolist()})
        if with_factors and i % 2 == 0:
            dur_f.append((0.8 + 0.4 * torch.rand(L, generator=g)).tolist())
            en_f.append((0.9 + 0.2 * torch.rand(L, generator=g)).tolist())
            pi_f.append(((torch.rand(L, generator=g) - 0.5) * 40.0).tolist() if pitch_transform == "add"
                        else (0.9 + 0.2 * torch.rand(L, generator=g)).tolist())
        else:
            dur_f.append(None)
            en_f.append(None)
            pi_f.append(None)
    return dict(
        sentences=sentences, file_names=[f"utt{i:04d}" for i in range(n)],
        speaker_ids=torch.randint(0, hp.n_speakers, (n,), generator=g).tolist(),
        dur_factors=dur_f, energy_factors=en_f, pitch_factors=[pitch_transform, pi_f], external_prosody=prosody,
        external_embeddings=torch.randn(n, getattr(hp, "external_emb_dim", 192), generator=g),
        external_accent_emb=torch.randn(n, hp.phoneme_encoder["hidden_embed_dim"], generator=g))
