"""Model-relevant hyper-parameters of the reference, as a plain namespace.

The reference's ``HyperParams`` class (``src/daft_exprt/hparams.py:19-161``) cannot be
constructed unaided (it asserts MFA files exist, ``hparams.py:218-229``) and never defines
``lambda_reversal`` although ``model.py:51`` reads it.  The hot path only reads the fields
below (SURVEY.md §5 "config"), so a namespace with the same attribute names is the whole
contract.  Values: ``hparams.py:41-46,72,106-127``.
"""
from types import SimpleNamespace
import copy


def english_symbols():
    """The reference's English symbol table (``symbols.py:16-40``, selected at ``hparams.py:178-191``): pad, EOS, whitespace,
    punctuation, then the stressed Arpabet inventory in alphabetical order of the base phone (vowels carry stress 0-2)."""
    vowels = {"AA", "AE", "AH", "AO", "AW", "AY", "EH", "ER", "EY", "IH", "IY", "OW", "OY", "UH", "UW"}
    consonants = {"B", "CH", "D", "DH", "F", "G", "HH", "JH", "K", "L", "M", "N", "NG", "P", "R", "S", "SH", "T", "TH", "V",
                  "W", "Y", "Z", "ZH"}
    phones = []
    for base in sorted(vowels | consonants):
        phones += [base + str(k) for k in range(3)] if base in vowels else [base]
    return list("_~ ,.!?") + phones


def default_hparams(**overrides):
    hp = SimpleNamespace(
        symbols=english_symbols(),
        n_mel_channels=80,
        n_symbols=76,
        n_speakers=2,
        lambda_reversal=1.0,
        post_mult_weight=1e-3,
        sampling_rate=22050,
        filter_length=1024,
        hop_length=256,
        centered=False,
        external_emb_dim=192,
        stats={
            "spk 0": {"pitch": {"mean": 5.0, "std": 0.25}, "energy": {"mean": 0.0, "std": 1.0}},
            "spk 1": {"pitch": {"mean": 5.3, "std": 0.30}, "energy": {"mean": 0.1, "std": 1.1}},
        },
        phoneme_encoder={
            "nb_blocks": 4, "hidden_embed_dim": 128, "attn_nb_heads": 2, "attn_dropout": 0.1,
            "conv_kernel": 3, "conv_channels": 1024, "conv_dropout": 0.1,
        },
        gaussian_upsampling_module={"conv_kernel": 3},
        frame_decoder={
            "nb_blocks": 4, "attn_nb_heads": 2, "attn_dropout": 0.1,
            "conv_kernel": 3, "conv_channels": 1024, "conv_dropout": 0.1,
        },
    )
    # training-side fields (BASELINE config 5): loss weights (hparams.py:72-88), optimiser (hparams.py:93-100)
    hp.mel_spec_weight = 1.0
    hp.energy_consistency_weight = 0.05
    hp.pitch_consistency_weight = 0.0
    hp.pitch_predictor_path = ""
    hp.adv_max_weight = 1e-2
    hp.warmup_steps = 10000
    hp.accumulation_steps = 1
    hp.betas = [0.9, 0.98]
    hp.epsilon = 1e-9
    hp.weight_decay = 1e-6
    hp.grad_clip_thresh = float("inf")
    hp.initial_learning_rate = 1e-4
    hp.max_learning_rate = 1e-3
    for k, v in overrides.items():
        setattr(hp, k, v)
    return hp


def clone_hparams(hp):
    """Deep copy (the reference's FrameDecoder mutates ``hparams.frame_decoder`` at init,
    ``model.py:534``; callers that build two models from one namespace want a copy)."""
    return copy.deepcopy(hp)


def accent_encoder_params(hp):
    """``AccentEncoder`` / ``StyleAdapter`` fall back to the phoneme-encoder sub-dict when
    ``hparams.accent_encoder`` is missing (``model.py:627-631,727-730``)."""
    return getattr(hp, "accent_encoder", None) or hp.phoneme_encoder
