"""TEST INFRASTRUCTURE ONLY — CPU restatement ("port") of Daft-Exprt's synthesis forward pass.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import this file, and only as the checker / the timed CPU baseline.  The
product (``ubisoft-laforge-daft-exprt_b200/``) never imports it and has no CPU fallback.

What it is: a functional fp32 PyTorch restatement of the math of the reference's
``DaftExprt.forward`` / ``DaftExprt.inference`` (``/root/reference/src/daft_exprt/model.py``),
written from the algorithm, operating directly on a ``state_dict`` in the
reference's checkpoint layout.  Each function cites the reference lines it restates.

Parity pinning: the reference ships **no** tests, golden vectors or fixtures, so the
pin is the live reference itself, imported in the build container: ``tests/test_oracle_vs_reference.py``
checks this port against it (integers bit-exact, floats ≤ 2e-5 max-abs) and
``tests/golden/make_golden.py`` freezes reference outputs into ``tests/golden/*.npz``, which travel to
the GPU box and are re-checked there against this port and against the CUDA path.
"""
import math
import torch
import torch.nn.functional as F

LN_EPS = 1e-5


# ----------------------------------------------------------------------------------------------
# small pieces
# ----------------------------------------------------------------------------------------------
def pad_mask(lengths, width=None):
    """True at padded positions.  Reference: ``get_mask_from_lengths`` (model.py:14-24), negated.
    The reference's width is ``max(lengths)``; callers' tensors have that width by construction."""
    if width is None:
        width = int(lengths.max())
    ids = torch.arange(width, dtype=torch.long)
    return ~(ids.unsqueeze(0) < lengths.unsqueeze(1))


_POS_CACHE = {}


def pos_table(dim, max_len=5000, timestep=10000.0):
    """Sin/cos table, fp32, same op sequence as model.py:125-130 so it is bit-identical."""
    key = (dim, max_len, timestep, str(torch.empty(0).device))     # per default device (bench.py also runs the port on the GPU)
    if key not in _POS_CACHE:
        pos = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
        div = torch.exp(torch.arange(0, dim, 2).float() * (-math.log(timestep) / dim))
        tab = torch.zeros(max_len, dim, dtype=torch.float32)
        tab[:, 0::2] = torch.sin(pos * div)
        tab[:, 1::2] = torch.cos(pos * div)
        _POS_CACHE[key] = tab
    return _POS_CACHE[key]


def pos_encoding(lengths, dim, width=None):
    """``PositionalEncoding.forward(lengths.unsqueeze(1))`` (model.py:132-150): row b gets table rows
    0..len_b-1, zeros after.  The gather index for (b, t) is simply t (integer contract)."""
    if width is None:
        width = int(lengths.max())
    tab = pos_table(dim)[:width]
    valid = ~pad_mask(lengths, width)
    return tab.unsqueeze(0) * valid.unsqueeze(2).to(tab.dtype)


def conv1d_same(x, w, b):
    """``ConvNorm1D`` (model.py:75-94): (B,L,Cin) → (B,L,Cout), zero padding (k-1)/2."""
    k = w.shape[2]
    return F.conv1d(x.transpose(1, 2), w, b, padding=(k - 1) // 2).transpose(1, 2)


def multi_head_attention(x, pad, w_in, b_in, w_out, b_out, n_heads):
    """Slow path of ``nn.MultiheadAttention`` as called at model.py:182-186 (q=k=v=x, key padding
    mask, eval mode).  scores = (q·d^-0.5)·kᵀ + mask; softmax; ·v; out-proj."""
    B, N, D = x.shape
    d = D // n_heads
    qkv = F.linear(x, w_in, b_in)
    q, k, v = qkv.split(D, dim=2)

    def heads(t):
        return t.reshape(B, N, n_heads, d).permute(0, 2, 1, 3).reshape(B * n_heads, N, d)

    q, k, v = heads(q), heads(k), heads(v)
    mask = torch.zeros(B, 1, 1, N, dtype=x.dtype).masked_fill(pad.view(B, 1, 1, N), float("-inf"))
    mask = mask.expand(B, n_heads, 1, N).reshape(B * n_heads, 1, N)
    scores = torch.baddbmm(mask, q * math.sqrt(1.0 / d), k.transpose(1, 2))
    attn = torch.softmax(scores, dim=-1)
    o = torch.bmm(attn, v).reshape(B, n_heads, N, d).permute(0, 2, 1, 3).reshape(B, N, D)
    return F.linear(o, w_out, b_out)


def fft_block(sd, prefix, x, film, pad, n_heads):
    """``FFTBlock.forward`` (model.py:249-259) = MultiHeadAttention.forward (171-193) +
    PositionWiseConvFF.forward (220-235) with the two pad masks.  Convs run on the padded layout
    (no mask between conv1 and conv2: hard part H1)."""
    a = prefix + ".attention."
    f = prefix + ".feed_forward."
    D = x.shape[2]
    o = multi_head_attention(x, pad, sd[a + "multi_head_attention.in_proj_weight"],
                             sd[a + "multi_head_attention.in_proj_bias"],
                             sd[a + "multi_head_attention.out_proj.weight"],
                             sd[a + "multi_head_attention.out_proj.bias"], n_heads)
    y = F.layer_norm(o + x, (D,), sd[a + "layer_norm.weight"], sd[a + "layer_norm.bias"], LN_EPS)
    y = y.masked_fill(pad.unsqueeze(2), 0)
    h = torch.relu(conv1d_same(y, sd[f + "convs.0.conv.weight"], sd[f + "convs.0.conv.bias"]))
    z = conv1d_same(h, sd[f + "convs.2.conv.weight"], sd[f + "convs.2.conv.bias"])
    z = F.layer_norm(z + y, (D,), sd[f + "layer_norm.weight"], sd[f + "layer_norm.bias"], LN_EPS)
    if film is not None:
        z = film[:, :D].unsqueeze(1) * z + film[:, D:].unsqueeze(1)
    return z.masked_fill(pad.unsqueeze(2), 0)


# ----------------------------------------------------------------------------------------------
# modules
# ----------------------------------------------------------------------------------------------
def _ae_params(hp):
    return getattr(hp, "accent_encoder", None) or hp.phoneme_encoder


def accent_encoder(sd, hp, frames_energy, frames_pitch, mel_specs, output_lengths):
    """``AccentEncoder.forward`` (model.py:679-716)."""
    ae = _ae_params(hp)
    D = ae["hidden_embed_dim"]
    T = mel_specs.shape[2]
    p = "accent_encoder."
    pos = pos_encoding(output_lengths, D, T)
    energy = conv1d_same(frames_energy.unsqueeze(2), sd[p + "energy_embedding.conv.weight"],
                         sd[p + "energy_embedding.conv.bias"])
    pitch = conv1d_same(frames_pitch.unsqueeze(2), sd[p + "pitch_embedding.conv.weight"],
                        sd[p + "pitch_embedding.conv.bias"])
    z = mel_specs.transpose(1, 2)
    for i in range(3):
        z = conv1d_same(z, sd[p + f"convs.{4 * i}.conv.weight"], sd[p + f"convs.{4 * i}.conv.bias"])
        z = torch.relu(z)
        z = F.layer_norm(z, (z.shape[2],), sd[p + f"convs.{4 * i + 2}.weight"],
                         sd[p + f"convs.{4 * i + 2}.bias"], LN_EPS)
    pad = pad_mask(output_lengths, T)
    x = (z + energy + pitch + pos).masked_fill(pad.unsqueeze(2), 0)
    for b in range(ae["nb_blocks"]):
        x = fft_block(sd, p + f"blocks.{b}", x, None, pad, ae["attn_nb_heads"])
    return torch.sum(x, dim=1) / output_lengths.unsqueeze(1)


def speaker_projection(sd, spk_embs):
    """model.py:904-905 / 1039-1040: L2-normalise (eps 1e-12) then Linear(192→128)."""
    e = F.normalize(spk_embs, p=2, dim=-1)
    return F.linear(e, sd["spk_projection.linear_layer.weight"], sd["spk_projection.linear_layer.bias"])


def speaker_classifier(sd, accent_emb):
    """``SpeakerClassifier.forward`` (model.py:829-830); the GRL is identity in forward."""
    p = "speaker_classifier.classifier."
    h = torch.relu(F.linear(accent_emb, sd[p + "1.linear_layer.weight"], sd[p + "1.linear_layer.bias"]))
    h = torch.relu(F.linear(h, sd[p + "3.linear_layer.weight"], sd[p + "3.linear_layer.bias"]))
    return F.linear(h, sd[p + "5.linear_layer.weight"], sd[p + "5.linear_layer.bias"])


def style_adapter(sd, hp, style_emb):
    """``StyleAdapter.forward`` (model.py:764-806) → {'phoneme_encoder','frame_decoder'}: (B,nb,2D)."""
    D = hp.phoneme_encoder["hidden_embed_dim"]
    nbs = [("phoneme_encoder", hp.phoneme_encoder["nb_blocks"]), ("frame_decoder", hp.frame_decoder["nb_blocks"])]
    g = F.linear(style_emb, sd["style_adapter.gammas_predictor.linear_layer.weight"],
                 sd["style_adapter.gammas_predictor.linear_layer.bias"])
    q = F.linear(style_emb, sd["style_adapter.betas_predictor.linear_layer.weight"],
                 sd["style_adapter.betas_predictor.linear_layer.bias"])
    post = sd.get("style_adapter.post_multipliers", None)
    B = style_emb.shape[0]
    out, col, blk = {}, 0, 0
    for name, nb in nbs:
        mg = g[:, col:col + nb * D].reshape(B, nb, D)
        mq = q[:, col:col + nb * D].reshape(B, nb, D)
        if post is not None:
            mg = post[0, blk:blk + nb].view(1, nb, 1) * mg + 1
            mq = post[1, blk:blk + nb].view(1, nb, 1) * mq
        else:
            mg = mg + 1
        out[name] = torch.cat((mg, mq), dim=2)
        col += nb * D
        blk += nb
    return out


def phoneme_encoder(sd, hp, symbols, film, input_lengths):
    """``PhonemeEncoder.forward`` (model.py:590-610)."""
    pe = hp.phoneme_encoder
    D = pe["hidden_embed_dim"]
    L = symbols.shape[1]
    x = F.embedding(symbols, sd["phoneme_encoder.symbols_embedding.weight"])
    pad = pad_mask(input_lengths, L)
    x = (x + pos_encoding(input_lengths, D, L)).masked_fill(pad.unsqueeze(2), 0)
    for b in range(pe["nb_blocks"]):
        fp = film[:, b, :] if film is not None else None
        x = fft_block(sd, f"phoneme_encoder.blocks.{b}", x, fp, pad, pe["attn_nb_heads"])
    return x


def length_regulator_integers(durations_int, input_lengths):
    """Integer side of ``GaussianUpsamplingModule.forward`` (model.py:485-497) and of
    ``inference`` (1096-1100): the quantities SURVEY.md §8a names the "length-regulator indices".
    All int64, bit-exact contract.
      cumsum[b,l]      inclusive scan of durations_int
      starts[b,l]      exclusive scan
      two_mu[b,l]      2·μ = d_l + 2·starts_l
      output_lengths   Σ_l d (no clamp here)
      t_max            max_b cumsum[b,-1]
      frame2symbol     (B,t_max) searchsorted(cumsum[b,:len_b], t, right=True); -1 for t ≥ out_len_b
    """
    d = durations_int.long()
    cs = torch.cumsum(d, dim=1)
    starts = cs - d
    two_mu = d + 2 * starts
    out_len = d.sum(dim=1)
    t_max = int(cs.max())
    B = d.shape[0]
    f2s = torch.full((B, t_max), -1, dtype=torch.long)
    t = torch.arange(t_max, dtype=torch.long)
    for b in range(B):
        n = int(input_lengths[b])
        idx = torch.searchsorted(cs[b, :n].contiguous(), t, right=True)
        ok = t < out_len[b]
        f2s[b, ok] = idx[ok]
    return {"cumsum": cs, "starts": starts, "two_mu": two_mu, "output_lengths": out_len,
            "t_max": t_max, "frame2symbol": f2s}


def gaussian_upsampling(sd, hp, x, durations_float, durations_int, energies, pitch, input_lengths,
                        t_max=None):
    """``GaussianUpsamplingModule.forward`` (model.py:417-510), film_params=None, summation path."""
    p = "gaussian_upsampling."
    dproj = conv1d_same(durations_float.unsqueeze(2), sd[p + "duration_projection.conv.weight"],
                        sd[p + "duration_projection.conv.bias"])
    eproj = conv1d_same(energies.unsqueeze(2), sd[p + "energy_projection.conv.weight"],
                        sd[p + "energy_projection.conv.bias"])
    pproj = conv1d_same(pitch.unsqueeze(2), sd[p + "pitch_projection.conv.weight"],
                        sd[p + "pitch_projection.conv.bias"])
    x = x + eproj + pproj
    r = F.softplus(F.linear(x + dproj, sd[p + "projection.0.linear_layer.weight"],
                            sd[p + "projection.0.linear_layer.bias"])).squeeze(2)
    pad = pad_mask(input_lengths, x.shape[1])
    sigma = torch.clamp(r.masked_fill(pad, 1), min=1e-3)
    means = durations_int.float() / 2
    cs = torch.cumsum(durations_int, dim=1)
    means[:, 1:] += cs[:, :-1]
    if t_max is None:
        t_max = int(cs.max())
    t = torch.arange(t_max, dtype=torch.float32) + 0.5
    mu = means.unsqueeze(-1)
    sg = sigma.unsqueeze(-1)
    # torch.distributions.Normal.log_prob: -((v-μ)²)/(2σ²) - log σ - log √(2π)
    logp = -((t - mu) ** 2) / (2 * sg ** 2) - sg.log() - math.log(math.sqrt(2 * math.pi))
    probs = torch.exp(logp).masked_fill(pad.unsqueeze(2), 0)
    w = probs / (probs.sum(dim=1, keepdim=True) + 1e-20)
    # same contraction as the reference (materialises (B,L,D,T), model.py:507) so that the CPU
    # baseline does the reference's work
    x_up = torch.sum(x.unsqueeze(-1) * w.unsqueeze(2), dim=1).permute(0, 2, 1)
    return x_up, w


def frame_decoder(sd, hp, x, film, output_lengths):
    """``FrameDecoder.forward`` (model.py:548-564)."""
    fd = hp.frame_decoder
    D = x.shape[2]
    T = x.shape[1]
    pad = pad_mask(output_lengths, T)
    x = (x + pos_encoding(output_lengths, D, T)).masked_fill(pad.unsqueeze(2), 0)
    for b in range(fd["nb_blocks"]):
        x = fft_block(sd, f"frame_decoder.blocks.{b}", x, film[:, b, :], pad, fd["attn_nb_heads"])
    mel = F.linear(x, sd["frame_decoder.projection.linear_layer.weight"],
                   sd["frame_decoder.projection.linear_layer.bias"])
    return mel.masked_fill(pad.unsqueeze(2), 0).transpose(1, 2)


# ----------------------------------------------------------------------------------------------
# whole passes
# ----------------------------------------------------------------------------------------------
@torch.no_grad()
def forward(sd, hp, inputs, external_accent_emb=None, external_spk_emb=None):
    """``DaftExprt.forward`` (model.py:889-948).  Same 5-tuple return."""
    if len(inputs) != 12:
        raise ValueError(f"inputs must have 12 elements (including spk_embs). Got {len(inputs)}.")
    (symbols, durations_float, durations_int, symbols_energy, symbols_pitch, input_lengths,
     frames_energy, frames_pitch, mel_specs, output_lengths, speaker_ids, spk_embs) = inputs
    if external_spk_emb is not None:
        spk_emb = external_spk_emb
    else:
        if spk_embs is None:
            raise ValueError("Speaker embeddings (spk_embs) required.")
        spk_emb = speaker_projection(sd, spk_embs)
    if external_accent_emb is not None:
        accent_emb = external_accent_emb
    else:
        accent_emb = accent_encoder(sd, hp, frames_energy, frames_pitch, mel_specs, output_lengths)
    speaker_preds = speaker_classifier(sd, accent_emb)
    film = style_adapter(sd, hp, accent_emb + spk_emb)
    enc = phoneme_encoder(sd, hp, symbols, film["phoneme_encoder"], input_lengths)
    x, weights = gaussian_upsampling(sd, hp, enc, durations_float, durations_int, symbols_energy,
                                     symbols_pitch, input_lengths)
    mel = frame_decoder(sd, hp, x, film["frame_decoder"], output_lengths)
    post = sd.get("style_adapter.post_multipliers", 1.0)
    film_params = [post, None, None, film["frame_decoder"]]
    encoder_preds = [durations_float, symbols_energy, symbols_pitch, input_lengths]
    decoder_preds = [mel, output_lengths]
    return speaker_preds, film_params, encoder_preds, decoder_preds, weights


def duration_to_integer(float_durations, hp):
    """``extract_features.duration_to_integer`` (extract_features.py:69-125), nb_samples=None.
    ``float_durations``: list of [begin, end] Python floats (seconds).  Returns list of ints; like the
    reference it may return one entry more or fewer than given (callers index-assign and fail)."""
    sr, flt, hop = hp.sampling_rate, hp.filter_length, hp.hop_length
    total = 0
    for b, e in float_durations:          # Python's sum(): starts from int 0, adds left to right
        total = total + (e - b)
    nb_samples = int(total * sr)
    nb_frames = 1 + int((nb_samples - flt) / hop)
    half = int(flt / 2)
    out, curr, i = [], 1, 0
    while curr <= nb_frames:
        if i >= len(float_durations):
            raise IndexError("pop from empty list")
        b, e = float_durations[i]
        i += 1
        if b == e:
            raise ValueError
        bs, es = int(b * sr), int(e * sr)
        # frames at sample index half + hop*k, k in [0, nb_frames), with bs < idx <= es
        n = 0
        for k in range(nb_frames):
            idx = half + hop * k
            if bs < idx <= es:
                n += 1
        out.append(n)
        curr += n
    remaining = len(float_durations) - i
    if hp.centered:
        edge = int(flt / 2 / hop)
        out[0] += edge
        if remaining != 0:
            out.append(edge)
        else:
            out[-1] += edge
    else:
        extra = int((flt - hop) / hop)
        left = extra // 2
        right = extra - left
        out[0] += left
        if remaining != 0:
            out.append(right)
        else:
            out[-1] += right
    return out


def get_int_durations(duration_preds, hp):
    """``DaftExprt.get_int_durations`` (model.py:950-973).  Mutates and returns ``duration_preds``."""
    dur_min = (hp.filter_length / hp.sampling_rate) / 2
    duration_preds[duration_preds < dur_min] = 0.
    B, L = duration_preds.shape
    out = torch.zeros(B, L, dtype=torch.long)
    for b in range(B):
        end_prev, idxs, fl = 0., [], []
        row = duration_preds[b].tolist()          # fp32 values widened to Python doubles, as .item()
        for l in range(L):
            d = row[l]
            if d != 0.:
                idxs.append(l)
                fl.append([end_prev, end_prev + d])
                end_prev += d
        ints = duration_to_integer(fl, hp)
        if len(ints) != len(idxs):
            raise RuntimeError(f"shape mismatch: value tensor of shape [{len(ints)}] cannot be broadcast "
                               f"to indexing result of shape [{len(idxs)}]")
        out[b, idxs] = torch.tensor(ints, dtype=torch.long)
    return duration_preds, out


def pitch_shift(pitch_preds, pitch_factors, hp, speaker_ids):
    """``DaftExprt.pitch_shift`` (model.py:975-994), in place."""
    zero = pitch_preds == 0.
    for b in range(pitch_preds.shape[0]):
        st = hp.stats[f"spk {int(speaker_ids[b])}"]["pitch"]
        mean, std = st["mean"], st["std"]
        pitch_preds[b] = torch.exp(std * pitch_preds[b] + mean)
        pitch_preds[b] += pitch_factors[b]
        pitch_preds[b] = (torch.log(pitch_preds[b]) - mean) / std
    pitch_preds[zero] = 0.
    return pitch_preds


def pitch_multiply(pitch_preds, pitch_factors):
    """``DaftExprt.pitch_multiply`` (model.py:996-1024), in place."""
    for b in range(pitch_preds.shape[0]):
        row = pitch_preds[b]
        zero = row == 0.
        mean = row[~zero].mean()
        dev = (row - mean) * pitch_factors[b]
        row += dev
        row[zero] = 0.
    return pitch_preds


@torch.no_grad()
def inference(sd, hp, inputs, pitch_transform, external_prosody=None, external_embeddings=None,
              external_accent_emb=None):
    """``DaftExprt.inference`` (model.py:1026-1114).  Same 3-tuple return.  Like the reference it
    mutates ``external_prosody['pitch_preds']`` in place (H8)."""
    symbols, dur_factors, energy_factors, pitch_factors, input_lengths, speaker_ids = inputs
    if external_embeddings is None:
        raise ValueError("external_embeddings required for inference.")
    spk_emb = speaker_projection(sd, external_embeddings)
    if external_accent_emb is None:
        raise ValueError("external_accent_emb required for inference.")
    film = style_adapter(sd, hp, external_accent_emb + spk_emb)
    enc = phoneme_encoder(sd, hp, symbols, film["phoneme_encoder"], input_lengths)
    if external_prosody is None:
        raise ValueError("external_prosody must be provided for inference.")
    duration_preds = external_prosody["duration_preds"] * dur_factors
    duration_preds, durations_int = get_int_durations(duration_preds, hp)
    energy_preds = external_prosody["energy_preds"] * energy_factors
    pitch_preds = external_prosody["pitch_preds"]
    energy_preds[durations_int == 0] = 0.
    pitch_preds[durations_int == 0] = 0.
    if pitch_transform == "add":
        pitch_preds = pitch_shift(pitch_preds, pitch_factors, hp, speaker_ids)
    elif pitch_transform == "multiply":
        pitch_preds = pitch_multiply(pitch_preds, pitch_factors)
    else:
        raise NotImplementedError
    x, weights = gaussian_upsampling(sd, hp, enc, duration_preds, durations_int, energy_preds,
                                     pitch_preds, input_lengths)
    output_lengths = torch.sum(durations_int, dim=1).long()
    output_lengths[output_lengths == 0] = 1
    assert int(output_lengths.max()) == x.shape[1]
    mel = frame_decoder(sd, hp, x, film["frame_decoder"], output_lengths)
    return ([duration_preds, durations_int, energy_preds, pitch_preds, input_lengths],
            [mel, output_lengths], weights)
