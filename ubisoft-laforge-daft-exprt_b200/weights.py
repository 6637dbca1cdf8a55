"""Checkpoint layout of the reference ``DaftExprt`` and synthetic random-init weights.

``state_dict_spec`` lists every tensor name and shape the reference's ``state_dict()`` holds
(184 tensors at default hparams; SURVEY.md §8b "Checkpoint"; reference ``model.py:838-856``).
``make_state_dict`` fills that layout with deterministic random values of xavier-like scale
(there is no network here for real checkpoints).  It does not use the reference's code: the
values come from one seeded ``torch.Generator`` walked over the sorted key list, so the same
weights can be rebuilt on any box and loaded into the reference model (strict) or ours.
"""
from collections import OrderedDict
import math
import torch

from .hparams import accent_encoder_params


def _fft_block_spec(prefix, D, C, k):
    a = prefix + ".attention."
    f = prefix + ".feed_forward."
    return [
        (a + "multi_head_attention.in_proj_weight", (3 * D, D)),
        (a + "multi_head_attention.in_proj_bias", (3 * D,)),
        (a + "multi_head_attention.out_proj.weight", (D, D)),
        (a + "multi_head_attention.out_proj.bias", (D,)),
        (a + "layer_norm.weight", (D,)),
        (a + "layer_norm.bias", (D,)),
        (f + "convs.0.conv.weight", (C, D, k)),
        (f + "convs.0.conv.bias", (C,)),
        (f + "convs.2.conv.weight", (D, C, k)),
        (f + "convs.2.conv.bias", (D,)),
        (f + "layer_norm.weight", (D,)),
        (f + "layer_norm.bias", (D,)),
    ]


def state_dict_spec(hp):
    """Ordered list of (name, shape) in the reference's registration order."""
    pe = hp.phoneme_encoder
    ae = accent_encoder_params(hp)
    fd = hp.frame_decoder
    D = pe["hidden_embed_dim"]
    Da, Ca, ka = ae["hidden_embed_dim"], ae["conv_channels"], ae["conv_kernel"]
    M = hp.n_mel_channels
    E = getattr(hp, "external_emb_dim", 192)
    kg = hp.gaussian_upsampling_module["conv_kernel"]
    spec = []
    # accent encoder (model.py:614-677)
    spec += [("accent_encoder.energy_embedding.conv.weight", (Da, 1, ka)),
             ("accent_encoder.energy_embedding.conv.bias", (Da,)),
             ("accent_encoder.pitch_embedding.conv.weight", (Da, 1, ka)),
             ("accent_encoder.pitch_embedding.conv.bias", (Da,))]
    chans = [(M, Ca), (Ca, Ca), (Ca, Da)]
    for i, (cin, cout) in enumerate(chans):
        spec += [(f"accent_encoder.convs.{4 * i}.conv.weight", (cout, cin, ka)),
                 (f"accent_encoder.convs.{4 * i}.conv.bias", (cout,)),
                 (f"accent_encoder.convs.{4 * i + 2}.weight", (cout,)),
                 (f"accent_encoder.convs.{4 * i + 2}.bias", (cout,))]
    for b in range(ae["nb_blocks"]):
        spec += _fft_block_spec(f"accent_encoder.blocks.{b}", Da, Ca, ka)
    # speaker classifier (model.py:809-830)
    spec += [("speaker_classifier.classifier.1.linear_layer.weight", (D, D)),
             ("speaker_classifier.classifier.1.linear_layer.bias", (D,)),
             ("speaker_classifier.classifier.3.linear_layer.weight", (D, D)),
             ("speaker_classifier.classifier.3.linear_layer.bias", (D,)),
             ("speaker_classifier.classifier.5.linear_layer.weight", (hp.n_speakers, D)),
             ("speaker_classifier.classifier.5.linear_layer.bias", (hp.n_speakers,))]
    # style adapter (model.py:719-762)
    nfilm = pe["nb_blocks"] * D + fd["nb_blocks"] * D
    if getattr(hp, "post_mult_weight", 0.0) != 0.0:
        spec += [("style_adapter.post_multipliers", (2, pe["nb_blocks"] + fd["nb_blocks"]))]
    spec += [("style_adapter.gammas_predictor.linear_layer.weight", (nfilm, Da)),
             ("style_adapter.gammas_predictor.linear_layer.bias", (nfilm,)),
             ("style_adapter.betas_predictor.linear_layer.weight", (nfilm, Da)),
             ("style_adapter.betas_predictor.linear_layer.bias", (nfilm,))]
    # phoneme encoder (model.py:567-588)
    spec += [("phoneme_encoder.symbols_embedding.weight", (hp.n_symbols, D))]
    for b in range(pe["nb_blocks"]):
        spec += _fft_block_spec(f"phoneme_encoder.blocks.{b}", D, pe["conv_channels"], pe["conv_kernel"])
    # gaussian upsampling (model.py:393-415)
    for n in ("duration", "energy", "pitch"):
        spec += [(f"gaussian_upsampling.{n}_projection.conv.weight", (D, 1, kg)),
                 (f"gaussian_upsampling.{n}_projection.conv.bias", (D,))]
    spec += [("gaussian_upsampling.projection.0.linear_layer.weight", (1, D)),
             ("gaussian_upsampling.projection.0.linear_layer.bias", (1,))]
    # frame decoder (model.py:513-546)
    Dd = getattr(hp, "frame_decoder_input_dim", D)
    for b in range(fd["nb_blocks"]):
        spec += _fft_block_spec(f"frame_decoder.blocks.{b}", Dd, fd["conv_channels"], fd["conv_kernel"])
    spec += [("frame_decoder.projection.linear_layer.weight", (M, Dd)),
             ("frame_decoder.projection.linear_layer.bias", (M,))]
    # speaker projection (model.py:855-856)
    spec += [("spk_projection.linear_layer.weight", (D, E)),
             ("spk_projection.linear_layer.bias", (D,))]
    return spec


def make_state_dict(hp, seed=1234, dtype=torch.float32):
    """Deterministic synthetic weights in the reference's checkpoint layout.

    Matrices/conv kernels: uniform(-a, a) with the xavier bound a = sqrt(6/(fan_in+fan_out));
    biases: uniform(-0.1, 0.1); LayerNorm weights: 1 + uniform(-0.1, 0.1) (so the affine part
    is exercised); post_multipliers: uniform(-0.6, 0.6).  One generator, keys in spec order.
    """
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    sd = OrderedDict()
    for name, shape in state_dict_spec(hp):
        if name.endswith("post_multipliers"):
            t = (torch.rand(shape, generator=g) * 2 - 1) * 0.6
        elif len(shape) == 1:
            t = (torch.rand(shape, generator=g) * 2 - 1) * 0.1
            is_ln = ("layer_norm.weight" in name) or (
                name.startswith("accent_encoder.convs.") and ".conv." not in name and name.endswith(".weight"))
            if is_ln:
                t = t + 1.0
        else:
            fan_out = shape[0] * (shape[2] if len(shape) == 3 else 1)
            fan_in = shape[1] * (shape[2] if len(shape) == 3 else 1)
            a = math.sqrt(6.0 / (fan_in + fan_out))
            t = (torch.rand(shape, generator=g) * 2 - 1) * a
        sd[name] = t.to(dtype).contiguous()
    return sd


def strip_module_prefix(state_dict):
    """Checkpoints saved from DDP carry a ``module.`` prefix; the reference's loaders strip it
    (``fine_tune.py:40``, ``scripts/synthesize.py:204``)."""
    return OrderedDict((k[7:] if k.startswith("module.") else k, v) for k, v in state_dict.items())
