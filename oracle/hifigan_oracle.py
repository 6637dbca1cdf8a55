"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's HiFi-GAN generator (mel → waveform).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU baseline may import this module; the product
(``ubisoft-laforge-daft-exprt_b200/vocoder.py``) never does.  SURVEY.md §8f row N4.

Restates ``/root/reference/src/daft_exprt/vocoder/hifigan.py``:
  * ``HiFiGANGenerator.forward``  (hifigan.py:159-176)  conv_pre → 4 × (leaky_relu → ConvTranspose1d → mean of three
    ResBlock1) → leaky_relu → conv_post → tanh; LRELU_SLOPE = 0.1 everywhere, including in front of conv_post;
  * ``ResBlock1.forward``         (hifigan.py:90-97)    x += conv2(leaky_relu(conv1_dilated(leaky_relu(x)))), 3 times;
  * ``remove_weight_norm``        (hifigan.py:178-184)  w = g · v / ‖v‖ over all dims but 0 (Conv1d: per output
    channel; ConvTranspose1d: per INPUT channel, because dim 0 of its weight is the input channel);
  * ``HiFiGanVocoder.infer``      (hifigan.py:211-226)  one utterance at a time on the cropped mel, clip to [-1, 1].

Parity pin: the reference ships no vocoder fixtures and its default checkpoint is a download (hifigan.py:24,43-68),
so the pin is the live reference class with seeded random weights (tests/test_oracle_vs_reference.py,
tests/golden/make_vocoder_golden.py → tests/golden/voc_*.npz).
"""
import math

import torch
import torch.nn.functional as F

LRELU_SLOPE = 0.1            # hifigan.py:21

DEFAULT_CONFIG = {           # hifigan.py:28-41
    "sampling_rate": 22050,
    "upsample_rates": [8, 8, 2, 2],
    "upsample_kernel_sizes": [16, 16, 4, 4],
    "upsample_initial_channel": 512,
    "resblock": "1",
    "resblock_kernel_sizes": [3, 7, 11],
    "resblock_dilation_sizes": [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    "model_in_dim": 80,
}


def conv_names(config=DEFAULT_CONFIG):
    """[(state_dict prefix, kind, Cin, Cout, kernel, stride or dilation)] in forward order of the parameters."""
    out = [("conv_pre", "conv", config["model_in_dim"], config["upsample_initial_channel"], 7, 1)]
    c0 = config["upsample_initial_channel"]
    for i, (u, k) in enumerate(zip(config["upsample_rates"], config["upsample_kernel_sizes"])):
        out.append((f"ups.{i}", "convT", c0 // 2 ** i, c0 // 2 ** (i + 1), k, u))
    for i in range(len(config["upsample_rates"])):
        ch = c0 // 2 ** (i + 1)
        for j, (k, dils) in enumerate(zip(config["resblock_kernel_sizes"], config["resblock_dilation_sizes"])):
            idx = i * len(config["resblock_kernel_sizes"]) + j
            for l, d in enumerate(dils):
                out.append((f"resblocks.{idx}.convs1.{l}", "conv", ch, ch, k, d))
            for l, _ in enumerate(dils):
                out.append((f"resblocks.{idx}.convs2.{l}", "conv", ch, ch, k, 1))
    out.append(("conv_post", "conv", ch, 1, 7, 1))
    return out


def make_state_dict(config=DEFAULT_CONFIG, seed=4321, weight_norm=True, gain=1.1):
    """Seeded random generator weights with O(1) signal gain per layer (the reference's own init, N(0, 0.01²) at
    hifigan.py:73-74, makes every activation vanish and would test nothing).  ``weight_norm=True`` produces the
    ``weight_g`` / ``weight_v`` checkpoint form the reference loads before ``remove_weight_norm``."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, kind, cin, cout, k, _ in conv_names(config):
        taps_in = k if kind == "conv" else 2          # a transposed conv with k = 2·stride sees 2 taps per output
        std = gain / math.sqrt(cin * taps_in)
        shape = (cout, cin, k) if kind == "conv" else (cin, cout, k)
        v = torch.randn(shape, generator=g) * std
        b = torch.randn(cout, generator=g) * 0.1
        if weight_norm:
            norm = v.flatten(1).norm(dim=1).view(-1, 1, 1)
            scale = 0.75 + 0.5 * torch.rand(shape[0], 1, 1, generator=g)
            sd[name + ".weight_g"] = norm * scale
            sd[name + ".weight_v"] = v
        else:
            sd[name + ".weight"] = v
        sd[name + ".bias"] = b
    return sd


def fold_weight_norm(sd):
    """``remove_weight_norm`` (hifigan.py:178-184): plain ``.weight`` tensors from ``weight_g`` / ``weight_v``."""
    out = {}
    for k, v in sd.items():
        if k.endswith(".weight_v"):
            p = k[: -len(".weight_v")]
            out[p + ".weight"] = torch._weight_norm(v, sd[p + ".weight_g"], 0)
        elif not k.endswith(".weight_g"):
            out[k] = v
    return out


def generator_forward(sd, mel, config=DEFAULT_CONFIG):
    """HiFiGANGenerator.forward (hifigan.py:159-176) on mel (B, 80, T) → (B, 1, T · prod(upsample_rates)); the whole
    padded tensor is processed, exactly as the reference module would."""
    w = fold_weight_norm(sd)
    nk = len(config["resblock_kernel_sizes"])
    x = F.conv1d(mel, w["conv_pre.weight"], w["conv_pre.bias"], padding=3)
    for i, (u, k) in enumerate(zip(config["upsample_rates"], config["upsample_kernel_sizes"])):
        x = F.leaky_relu(x, LRELU_SLOPE)
        x = F.conv_transpose1d(x, w[f"ups.{i}.weight"], w[f"ups.{i}.bias"], stride=u, padding=(k - u) // 2)
        xs = None
        for j, (rk, dils) in enumerate(zip(config["resblock_kernel_sizes"], config["resblock_dilation_sizes"])):
            p = f"resblocks.{i * nk + j}"
            y = x
            for l, d in enumerate(dils):                                   # ResBlock1.forward, hifigan.py:90-97
                xt = F.leaky_relu(y, LRELU_SLOPE)
                xt = F.conv1d(xt, w[f"{p}.convs1.{l}.weight"], w[f"{p}.convs1.{l}.bias"], dilation=d,
                              padding=int((rk * d - d) / 2))
                xt = F.leaky_relu(xt, LRELU_SLOPE)
                xt = F.conv1d(xt, w[f"{p}.convs2.{l}.weight"], w[f"{p}.convs2.{l}.bias"], padding=int((rk - 1) / 2))
                y = xt + y
            xs = y if xs is None else xs + y
        x = xs / nk
    x = F.leaky_relu(x, LRELU_SLOPE)
    x = F.conv1d(x, w["conv_post.weight"], w["conv_post.bias"], padding=3)
    return torch.tanh(x)


def infer(sd, mel, config=DEFAULT_CONFIG):
    """HiFiGanVocoder.infer (hifigan.py:211-226): one (80, T) mel → clipped waveform (T · 256,)."""
    with torch.no_grad():
        audio = generator_forward(sd, mel.float().unsqueeze(0), config).squeeze()
    return torch.clamp(audio, -1.0, 1.0)


def infer_batch(sd, mels, lengths, config=DEFAULT_CONFIG):
    """What a synthesis driver does with a padded batch (generate.py vocoder loop): every utterance is cropped to
    its own length and vocoded alone.  Returns (B, T_max · hop) with zeros behind each utterance."""
    hop = 1
    for u in config["upsample_rates"]:
        hop *= u
    B, _, T = mels.shape
    out = torch.zeros(B, T * hop)
    for b in range(B):
        n = int(lengths[b])
        if n > 0:
            out[b, : n * hop] = infer(sd, mels[b, :, :n], config)
    return out
