"""ctypes binding of libdaft_b200.so (include/daft_b200.h).  Plain pointers and sizes only.

The product has NO CPU fallback: if the shared library is missing or a call fails, this raises.
"""
import ctypes as C
import os

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libdaft_b200.so")

MODE_F32, MODE_BF16 = 0, 1
IO_PACKED_FRAMES, IO_PACKED_MEL_OUT = 1, 2
GAP, HALO, TAIL = 3, 8, 264
D, HEADS, DH, CCH, TAPS, MEL, MEL_PAD, SPK = 128, 2, 64, 1024, 3, 80, 96, 192
MAX_BLOCKS = 4          # FFT blocks per stack: DxModelWeights holds 4 slots (1..4 used)

_p = C.c_void_p


class DxBlockWeights(C.Structure):
    _fields_ = [(n, _p) for n in (
        "w_qkv", "b_qkv", "w_out", "b_out", "ln1_g", "ln1_b",
        "w_c1", "b_c1", "w_c2", "b_c2", "ln2_g", "ln2_b",
        "w_qkv_bf", "w_out_bf", "w_c1_bf", "w_c2_bf", "w_ffn1_bf", "w_ffn2_bf")]


class DxModelWeights(C.Structure):
    _fields_ = (
        [(n, _p) for n in ("ae_energy_w", "ae_energy_b", "ae_pitch_w", "ae_pitch_b",
                           "ae_c0_w", "ae_c0_b", "ae_ln0_g", "ae_ln0_b",
                           "ae_c1_w", "ae_c1_b", "ae_ln1_g", "ae_ln1_b",
                           "ae_c2_w", "ae_c2_b", "ae_ln2_g", "ae_ln2_b",
                           "ae_c0_w_bf", "ae_c1_w_bf", "ae_c2_w_bf")]
        + [("ae_blocks", DxBlockWeights * 4)]
        + [(n, _p) for n in ("spk_w", "spk_b", "cls_w1", "cls_b1", "cls_w2", "cls_b2", "cls_w3", "cls_b3",
                             "film_g_w", "film_g_b", "film_b_w", "film_b_b", "post_mult",
                             "symbols_embedding")]
        + [("pe_blocks", DxBlockWeights * 4)]
        + [(n, _p) for n in ("gu_dur_w", "gu_dur_b", "gu_energy_w", "gu_energy_b", "gu_pitch_w", "gu_pitch_b",
                             "gu_range_w", "gu_range_b")]
        + [("fd_blocks", DxBlockWeights * 4)]
        + [(n, _p) for n in ("mel_w", "mel_b", "mel_w_bf", "pos_table")]
        + [("n_symbols", C.c_int32), ("n_speakers", C.c_int32), ("pos_rows", C.c_int32), ("pad_", C.c_int32)]
        + [("n_ae_blocks", C.c_int32), ("n_pe_blocks", C.c_int32), ("n_fd_blocks", C.c_int32), ("pad2_", C.c_int32)]
    )


class DxForwardArgs(C.Structure):
    _fields_ = (
        [("B", C.c_int32), ("L_max", C.c_int32), ("T_max", C.c_int32), ("mode", C.c_int32)]
        + [(n, _p) for n in ("symbols", "durations_float", "durations_int", "symbols_energy", "symbols_pitch",
                             "input_lengths", "frames_energy", "frames_pitch", "mel_specs", "output_lengths",
                             "spk_embs", "accent_emb_in", "spk_emb_in",
                             "accent_emb", "speaker_preds", "film", "enc_out", "mel_out", "weights_out",
                             "output_lengths_out", "workspace")]
        + [("workspace_bytes", C.c_int64), ("io_flags", C.c_int32), ("pad_", C.c_int32)]
    )


TRAIN_FORWARD_ONLY = 1
TRAIN_TC_GEMM = 2


class DxTrainArgs(C.Structure):
    _fields_ = (
        [("B", C.c_int32), ("L_max", C.c_int32), ("T_max", C.c_int32), ("flags", C.c_int32)]
        + [(n, _p) for n in ("symbols", "durations_float", "durations_int", "symbols_energy", "symbols_pitch", "input_lengths",
                             "frames_energy", "frames_pitch", "mel_specs", "output_lengths", "speaker_ids", "spk_embs")]
        + [(n, C.c_float) for n in ("mel_spec_weight", "post_mult_weight", "energy_consistency_weight", "speaker_weight",
                                    "lambda_reversal", "loss_scale")]
        + [("losses", _p), ("mel_out", _p), ("workspace", _p), ("workspace_bytes", C.c_int64)]
        + [("attn_dropout", C.c_float * 3), ("conv_dropout", C.c_float * 3), ("dropout_seed", C.c_uint64)]
    )


VC_GAP, VC_HALO, VC_RESIDUAL_F32 = 4, 32, 1


class DxVcConv(C.Structure):
    _fields_ = [("w", _p), ("w_bf", _p), ("bias", _p),
                ("taps", C.c_int32), ("dil", C.c_int32), ("K", C.c_int32), ("N", C.c_int32)]


class DxVocoderWeights(C.Structure):
    _fields_ = [("n_stages", C.c_int32), ("n_kernels", C.c_int32), ("n_layers", C.c_int32), ("mel_channels", C.c_int32),
                ("up", C.c_int32 * 4), ("channels", C.c_int32 * 4),
                ("pre", DxVcConv), ("ups", DxVcConv * 4), ("rb", (((DxVcConv * 2) * 3) * 3) * 4), ("post", DxVcConv)]


class DxVocoderArgs(C.Structure):
    _fields_ = [("B", C.c_int32), ("T_max", C.c_int32), ("mode", C.c_int32), ("flags", C.c_int32),
                ("frame_cap", C.c_int64), ("mel", _p), ("lengths", _p), ("audio", _p), ("workspace", _p),
                ("workspace_bytes", C.c_int64)]


# name -> (restype, argtypes); must list every function include/daft_b200.h declares
SIGNATURES = {
    "dx_version": (C.c_int, []),
    "dx_last_error_string": (C.c_char_p, []),
    "dx_launch_count": (C.c_int64, []),
    "dx_set_debug": (C.c_int, [C.c_int, C.c_int]),
    "dx_set_trace_buffer": (C.c_int, [_p, C.c_int64]),
    "dx_profile_enable": (C.c_int, [C.c_int]),
    "dx_profile_report": (C.c_int, [C.c_char_p, C.c_int64]),
    "dx_row_cap": (C.c_int64, [C.c_int, C.c_int]),
    "dx_build_rows": (C.c_int, [_p, C.c_int, C.c_int, _p, _p, _p, C.c_int64, _p]),
    "dx_packed_weight_elems": (C.c_int64, [C.c_int] * 5),
    "dx_pack_weight_bf16": (C.c_int, [_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _p, _p]),
    "dx_forward_workspace_bytes": (C.c_int64, [C.c_int] * 4),
    "dx_forward": (C.c_int, [C.POINTER(DxModelWeights), C.POINTER(DxForwardArgs), _p]),
    "dx_train_workspace_bytes": (C.c_int64, [C.c_int] * 3),
    "dx_dropout_mask": (C.c_int, [C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, _p, _p]),
    "dx_train_step": (C.c_int, [C.POINTER(DxModelWeights), C.POINTER(DxModelWeights), C.POINTER(DxTrainArgs), _p]),
    "dx_adam_step": (C.c_int, [_p, _p, _p, _p, C.c_int64] + [C.c_float] * 5 + [C.c_int, C.c_float, _p]),
    "dx_accent_encoder": (C.c_int, [C.POINTER(DxModelWeights), C.c_int, C.c_int, C.c_int, _p, _p, _p, _p, _p, _p,
                                    C.c_int64, _p]),
    "dx_style": (C.c_int, [C.POINTER(DxModelWeights), C.c_int, _p, _p, _p, _p, _p, _p]),
    "dx_length_regulator_integers": (C.c_int, [_p, _p, C.c_int, C.c_int, C.c_int, _p, _p, _p, _p, _p]),
    "dx_int_durations": (C.c_int, [_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _p, _p, _p]),
    "dx_prosody_transform": (C.c_int, [_p, _p, _p, _p, _p, _p, _p, C.c_int, C.c_int, C.c_int, _p]),
    "dx_tc_gemm_test": (C.c_int, [_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, _p, C.c_int, C.c_int, _p, _p]),
    "dx_vocoder_frame_cap": (C.c_int64, [C.c_int, C.c_int64]),
    "dx_vocoder_conv_tiling": (C.c_int, [C.c_int] * 4 + [C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "dx_vocoder_workspace_bytes": (C.c_int64, [C.POINTER(DxVocoderWeights), C.c_int, C.c_int64, C.c_int]),
    "dx_vocoder_forward": (C.c_int, [C.POINTER(DxVocoderWeights), C.POINTER(DxVocoderArgs), _p]),
    "dx_host_collate": (C.c_int, [C.c_int] * 5 + [_p, _p, _p, _p, C.c_int]),
    "dx_host_collate_packed": (C.c_int, [C.c_int] * 5 + [_p, _p, _p, _p, C.c_int]),
    "dx_test_attention": (C.c_int, [C.c_int, _p, _p, C.c_int64, _p, C.c_int, C.c_int, _p, C.c_int64, _p]),
    "dx_attention_static_max_rows": (C.c_int64, []),
    "dx_attention_static": (C.c_int, [C.c_int, _p, _p, C.c_int64, _p, C.c_int, C.c_int, _p]),
}

_LIB = None


class DxError(RuntimeError):
    pass


def load():
    """Load libdaft_b200.so (built by build.py / __graft_entry__.build()).  Raises if it is missing."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.isfile(LIB_PATH):
        raise DxError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                      f"(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)     # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    # bring-up switches, e.g. DX_DEBUG="0=1,1=1" (all 0 in production)
    for kv in filter(None, os.environ.get("DX_DEBUG", "").split(",")):
        k, v = kv.split("=")
        lib.dx_set_debug(int(k), int(v))
    _LIB = lib
    return lib


def check(status, what=""):
    if status != 0:
        msg = load().dx_last_error_string()
        raise DxError(f"{what} failed (status {status}): {msg.decode() if msg else ''}")


def ptr(t):
    """Device/host pointer of a tensor (None → NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    """The current CUDA stream of ``device`` (default: the current device) as the C ABI's ``void* stream``."""
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
