"""CPU: the HiFi-GAN oracle against the golden vectors frozen from the live reference generator, against the live
reference itself (build container only), and the host logic of the product mirror (checkpoint folding, polyphase
form of the transposed convolutions, config gate, loud failure without a GPU)."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import _cases as C
from oracle import hifigan_oracle as H
import make_vocoder_golden as VG
from daft_exprt_b200 import _lib
from daft_exprt_b200 import vocoder as V


@pytest.mark.parametrize("name", sorted(VG.VOCODER_CASES))
def test_oracle_matches_golden(name):
    g = C.load_golden(name)
    sd = H.make_state_dict(seed=VG.WEIGHT_SEED)
    mels, lengths = VG.vocoder_case(**VG.VOCODER_CASES[name])
    assert str(g["weights_digest"]) == VG.digest([sd[k] for k in sorted(sd)]), "weight generator drifted"
    assert str(g["inputs_digest"]) == VG.digest([mels, lengths]), "input generator drifted"
    audio = H.infer_batch(sd, mels, lengths).numpy()
    assert audio.shape == g["audio"].shape
    assert np.abs(audio - g["audio"]).max() <= 1e-6
    assert np.abs(g["audio"]).std() > 0.1            # the case is not saturated / vanishing


@pytest.mark.skipif(not os.path.isfile(VG.REF_FILE), reason="reference only exists in the build container")
def test_oracle_is_bit_identical_to_live_reference():
    sd = H.make_state_dict(seed=77)
    gen = VG.reference_generator(sd)
    mel = torch.randn(2, 80, 11, generator=torch.Generator().manual_seed(5))
    with torch.no_grad():
        ref = gen(mel)
        ours = H.generator_forward(sd, mel)
    assert torch.equal(ref, ours)
    # folded weights equal the reference's after remove_weight_norm
    folded = H.fold_weight_norm(sd)
    for k, v in gen.state_dict().items():
        assert torch.equal(v, folded[k]), k


def test_padded_batch_is_not_the_cropped_result():
    """Why the product masks per utterance: the reference module run on a zero-padded batch gives different samples
    near an utterance's end than the cropped call its driver makes (conv_pre's bias is live on the padded frames)."""
    sd = H.make_state_dict(seed=VG.WEIGHT_SEED)
    mels, lengths = VG.vocoder_case(lengths=[9, 5], seed=3)
    with torch.no_grad():
        padded = H.generator_forward(sd, mels)[1, 0, : 5 * 256]
    cropped = H.infer(sd, mels[1, :, :5])
    assert (padded - cropped).abs().max() > 1e-3


@pytest.mark.parametrize("u", [2, 8])
def test_polyphase_weight_equals_conv_transpose(u):
    g = torch.Generator().manual_seed(u)
    w, x, b = torch.randn(16, 32, 2 * u, generator=g), torch.randn(1, 16, 13, generator=g), torch.randn(32, generator=g)
    ref = F.conv_transpose1d(x, w, b, stride=u, padding=u // 2)[0]              # (32, 13 u)
    pw = V.polyphase_weight(w, u)
    assert pw.shape == (3, 16, u * 32)
    xp = F.pad(x[0].t(), (0, 0, 1, 1))
    out = sum(xp[t : t + 13] @ pw[t] for t in range(3)).reshape(13 * u, 32) + b
    assert (out.t() - ref).abs().max() <= 1e-5


def test_generator_mirror_checkpoint_forms_and_cpu_failure():
    sd = H.make_state_dict(seed=VG.WEIGHT_SEED)
    folded = H.fold_weight_norm(sd)
    gen = V.HiFiGANGenerator(V.DEFAULT_CONFIG)
    gen.load_state_dict(sd)                                  # weight_g / weight_v form
    got = gen.state_dict()
    assert sorted(got) == sorted(folded)
    assert all(torch.equal(got[k], folded[k]) for k in folded)
    gen2 = V.HiFiGANGenerator(V.DEFAULT_CONFIG)
    gen2.load_state_dict({"module." + k: v for k, v in folded.items()})   # plain form, DDP prefix
    assert all(torch.equal(gen2.state_dict()[k], folded[k]) for k in folded)
    with pytest.raises(RuntimeError):
        gen2.load_state_dict({k: v for k, v in folded.items() if k != "conv_post.bias"})
    if not torch.cuda.is_available():
        with pytest.raises(_lib.DxError):
            gen(torch.zeros(1, 80, 4))                      # no CPU fallback
    with pytest.raises(FileNotFoundError):
        V.HiFiGanVocoder(checkpoint_path=None)               # the reference would download here


def test_config_gate():
    bad = dict(V.DEFAULT_CONFIG, resblock='2')
    with pytest.raises(NotImplementedError):
        V.HiFiGANGenerator(bad)
    bad = dict(V.DEFAULT_CONFIG, upsample_kernel_sizes=[16, 16, 4, 3])
    with pytest.raises(NotImplementedError):
        V.HiFiGANGenerator(bad)


def test_vocoder_host_queries():
    lib = _lib.load()
    import ctypes
    assert lib.dx_vocoder_frame_cap(3, 100) == 128 and lib.dx_vocoder_frame_cap(3, 113) == 256
    nt, ks = ctypes.c_int(), ctypes.c_int()
    assert lib.dx_vocoder_conv_tiling(11, 5, 256, 256, ctypes.byref(nt), ctypes.byref(ks)) == 0
    assert nt.value == 128 and ks.value == 16
    assert lib.dx_vocoder_conv_tiling(3, 1, 32, 32, ctypes.byref(nt), ctypes.byref(ks)) == 0
    assert nt.value == 32 and ks.value == 32
    assert ctypes.sizeof(_lib.DxVcConv) == 40
    assert ctypes.sizeof(_lib.DxVocoderWeights) == 48 + 40 * (1 + 4 + 72 + 1)


def test_product_synthetic_checkpoint_equals_the_oracles():
    a, b = V.synthetic_state_dict(seed=5), H.make_state_dict(seed=5)
    assert list(a) == list(b) and all(torch.equal(a[k], b[k]) for k in a)
    assert V.flops_per_frame() == 614_105_088


def test_vocoder_cabi_refuses_bad_arguments_without_a_gpu():
    """Argument validation of dx_vocoder_forward / dx_vocoder_workspace_bytes happens before any CUDA call."""
    import ctypes
    lib = _lib.load()
    W = _lib.DxVocoderWeights()
    assert lib.dx_vocoder_workspace_bytes(ctypes.byref(W), 1, 128, _lib.MODE_BF16) == -1         # 0 stages
    assert b"stages" in lib.dx_last_error_string()
    W.n_stages, W.n_kernels, W.n_layers, W.mel_channels = 4, 3, 3, 80
    for i, (u, c) in enumerate(zip((8, 8, 2, 2), (256, 128, 64, 32))):
        W.up[i], W.channels[i] = u, c
    W.pre.N = 512
    assert lib.dx_vocoder_workspace_bytes(ctypes.byref(W), 1, 100, _lib.MODE_BF16) == -1         # frame_cap % 128
    n_bf = lib.dx_vocoder_workspace_bytes(ctypes.byref(W), 2, 256, _lib.MODE_BF16)
    n_f32 = lib.dx_vocoder_workspace_bytes(ctypes.byref(W), 2, 256, _lib.MODE_F32)
    assert 0 < n_bf < n_f32                                                                       # bf16 operands are half the size
    assert 256 * 150_000 < n_bf < 256 * 180_000                                                   # ~160 KB per frame
    W.up[2] = 3
    assert lib.dx_vocoder_workspace_bytes(ctypes.byref(W), 2, 256, _lib.MODE_BF16) == -1         # not a power of two
    W.up[2] = 2
    a = _lib.DxVocoderArgs()
    assert lib.dx_vocoder_forward(ctypes.byref(W), ctypes.byref(a), None) != 0                    # NULL tensors
    assert b"NULL" in lib.dx_last_error_string()
