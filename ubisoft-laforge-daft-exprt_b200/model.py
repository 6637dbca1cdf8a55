"""Drop-in mirror of the reference's ``daft_exprt.model.DaftExprt`` (reference ``model.py:833-1114``).

Same constructor, same ``state_dict`` layout (184 tensors), same ``forward`` /
``inference`` / ``parse_batch`` / ``accent_encoder(...)`` signatures, return tuples and exception
types.  The arithmetic runs in libdaft_b200.so (hand-written sm_100a CUDA behind the C ABI of
``include/daft_b200.h``); PyTorch only owns device memory and the stream.  There is no CPU
fallback: without the shared library or a CUDA device the calls raise.

``mode``: ``"bf16"`` (default; tcgen05 tensor-core kernels, fp32 accumulation/LayerNorm/softmax)
or ``"fp32"`` (SIMT fp32 kernels; the 1e-3 parity mode).
"""
import ctypes as C
import weakref

import torch
from torch import nn

from . import _lib
from .hparams import accent_encoder_params
from .weights import make_state_dict, state_dict_spec, strip_module_prefix

_MODES = {"fp32": _lib.MODE_F32, "f32": _lib.MODE_F32, "bf16": _lib.MODE_BF16}


class _Holder(nn.Module):
    """Parameter container: reproduces the reference's module tree so state_dict keys match."""


class _AccentEncoderHandle(_Holder):
    """``model.accent_encoder(frames_energy, frames_pitch, mel_specs, output_lengths)`` is called
    directly by the reference's synthesis script (``scripts/synthesize.py:438``)."""

    def __init__(self, owner):
        super().__init__()
        object.__setattr__(self, "_owner", weakref.ref(owner))

    def forward(self, frames_energy, frames_pitch, mel_specs, output_lengths):
        return self._owner()._accent_encoder(frames_energy, frames_pitch, mel_specs, output_lengths)


def _pos_table(dim, max_len=5000, timestep=10000.0):
    """Same torch op sequence as the reference (``model.py:125-130``) → bit-identical fp32 table."""
    import numpy as np
    pos = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
    div = torch.exp(torch.arange(0, dim, 2).float() * (-np.log(timestep) / dim))
    tab = torch.FloatTensor(max_len, dim).zero_()
    tab[:, 0::2] = torch.sin(pos * div)
    tab[:, 1::2] = torch.cos(pos * div)
    return tab


class DaftExprt(nn.Module):
    def __init__(self, hparams, is_training=True, mode="bf16", init_seed=1234):
        super().__init__()
        self._check_hparams(hparams)
        self.n_speakers = hparams.n_speakers
        self.hidden_embed_dim = hparams.phoneme_encoder["hidden_embed_dim"]
        self.compute_mode = mode
        if mode not in _MODES:
            raise ValueError(f"mode must be one of {sorted(_MODES)}")
        object.__setattr__(self, "_hp", hparams)
        # parameter tree with the reference's names / shapes / registration order
        init = make_state_dict(hparams, init_seed)
        for name, _ in state_dict_spec(hparams):
            parts = name.split(".")
            node = self
            for i, part in enumerate(parts[:-1]):
                child = node._modules.get(part)
                if child is None:
                    child = _AccentEncoderHandle(self) if (i == 0 and part == "accent_encoder") else _Holder()
                    node.add_module(part, child)
                node = child
            node.register_parameter(parts[-1], nn.Parameter(init[name].clone()))
        if getattr(hparams, "post_mult_weight", 0.0) == 0.0:
            self.style_adapter.post_multipliers = 1.0          # as the reference (model.py:760-761)
        # like the reference's FrameDecoder.__init__ (model.py:534)
        hparams.frame_decoder["hidden_embed_dim"] = self.hidden_embed_dim
        self._prepared = None        # (key, DxModelWeights, keep-alive tensors)
        self._param_list = None
        self.streams = 1             # utterance groups run concurrently per call (see _run)
        self._side_streams = []
        self._workspaces = {}
        self._stats_table = None

    # ------------------------------------------------------------------------------------------
    @staticmethod
    def _check_hparams(hp):
        pe, fd = hp.phoneme_encoder, hp.frame_decoder
        ae = accent_encoder_params(hp)
        ok = (pe["hidden_embed_dim"] == _lib.D and ae["hidden_embed_dim"] == _lib.D
              and pe["attn_nb_heads"] == _lib.HEADS and fd["attn_nb_heads"] == _lib.HEADS and ae["attn_nb_heads"] == _lib.HEADS
              and pe["conv_channels"] == _lib.CCH and fd["conv_channels"] == _lib.CCH and ae["conv_channels"] == _lib.CCH
              and pe["conv_kernel"] == 3 and fd["conv_kernel"] == 3 and ae["conv_kernel"] == 3
              and all(1 <= x["nb_blocks"] <= _lib.MAX_BLOCKS for x in (pe, fd, ae))
              and hp.gaussian_upsampling_module["conv_kernel"] == 3
              and hp.n_mel_channels == _lib.MEL and getattr(hp, "external_emb_dim", 192) == _lib.SPK
              and not hp.gaussian_upsampling_module.get("use_concatenation", False)
              and not hasattr(hp, "frame_decoder_input_dim"))
        if not ok:
            raise NotImplementedError(
                "the B200 kernels are specialised for the reference's default layer shapes "
                "(hparams.py:106-127: D=128, 2 heads, 1024 conv channels, k=3, 80 mels, 192-d ECAPA) with 1 to 4 FFT blocks per "
                "stack (phoneme_encoder / frame_decoder / accent_encoder 'nb_blocks' may differ from each other)")

    # invalidate packed weights whenever parameters may have changed / moved
    def _apply(self, fn, *a, **k):
        self._prepared = None
        self._param_list = None
        self._workspaces = {}
        self._stats_table = None
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, state_dict, strict=True, **kw):
        self._prepared = None
        self._param_list = None
        return super().load_state_dict(strip_module_prefix(state_dict), strict=strict, **kw)

    def set_mode(self, mode):
        if mode not in _MODES:
            raise ValueError(f"mode must be one of {sorted(_MODES)}")
        self.compute_mode = mode
        return self

    @property
    def device(self):
        return self.spk_projection.linear_layer.weight.device

    # ------------------------------------------------------------------------------------------
    # weight preparation: GEMM layouts (fp32) and tile-packed bf16 for the tensor-core kernels
    # ------------------------------------------------------------------------------------------
    def _param_versions(self):
        """Identity and version counter of every parameter, the key of the packed-weight cache.  The MODULE tree is fixed
        after construction, so the (module's parameter dict, name) slots are collected once: walking the tree with
        ``self.parameters()`` on every call cost a third of the host time of a forward (``named_modules`` recursion over 184
        parameters).  Reading through the dicts still sees a parameter OBJECT that was replaced (``m.x.weight =
        nn.Parameter(...)``) as well as in-place updates (optimizer steps, ``load_state_dict``: version bumps)."""
        slots = self._param_list
        if slots is None:
            slots = self._param_list = [(m._parameters, n) for m in self.modules() for n in m._parameters]
        return tuple((id(d[n]), d[n]._version) for d, n in slots)

    def _prepare(self):
        dev = self.device
        if dev.type != "cuda":
            raise _lib.DxError("DaftExprt (B200) needs its parameters on a CUDA device; there is no CPU fallback")
        mode = _MODES[self.compute_mode]
        key = (mode, str(dev), self._param_versions())
        if self._prepared is not None and self._prepared[0] == key:
            return self._prepared[1]
        lib = _lib.load()
        keep = []
        sd = {k: v.detach() for k, v in self.state_dict().items()}

        def f32(t):
            t = t.to(device=dev, dtype=torch.float32).contiguous()
            keep.append(t)
            return t

        def conv_w(name):     # torch conv weight [Cout][Cin][k] → GEMM layout [k][Cin][Cout]
            return f32(sd[name].permute(2, 1, 0))

        def lin_w(name):      # torch linear weight [out][in] → [1][in][out]
            return f32(sd[name].t().unsqueeze(0))

        def packed(w_gemm, nt, ks, k_pad=None):
            """w_gemm: [taps][K][N] fp32 on device → bf16 tile-packed (dx_pack_weight_bf16)."""
            if mode != _lib.MODE_BF16:
                return None
            taps, K, N = w_gemm.shape
            if k_pad is not None and k_pad != K:
                w_gemm = torch.cat([w_gemm, w_gemm.new_zeros(taps, k_pad - K, N)], dim=1).contiguous()
                keep.append(w_gemm)
                K = k_pad
            n = lib.dx_packed_weight_elems(taps, K, N, nt, ks)
            out = torch.empty(n, dtype=torch.bfloat16, device=dev)
            with torch.cuda.device(dev):
                _lib.check(lib.dx_pack_weight_bf16(_lib.ptr(w_gemm), taps, K, N, nt, ks, _lib.ptr(out), _lib.stream_ptr(dev)),
                           "dx_pack_weight_bf16")
            keep.append(out)
            return out

        W = _lib.DxModelWeights()

        def setp(struct, field, t):
            setattr(struct, field, None if t is None else t.data_ptr())

        def fill_block(bw, prefix):
            a, f = prefix + ".attention.", prefix + ".feed_forward."
            w_qkv = lin_w(a + "multi_head_attention.in_proj_weight")
            w_out = lin_w(a + "multi_head_attention.out_proj.weight")
            w_c1 = conv_w(f + "convs.0.conv.weight")
            w_c2 = conv_w(f + "convs.2.conv.weight")
            setp(bw, "w_qkv", w_qkv); setp(bw, "b_qkv", f32(sd[a + "multi_head_attention.in_proj_bias"]))
            setp(bw, "w_out", w_out); setp(bw, "b_out", f32(sd[a + "multi_head_attention.out_proj.bias"]))
            setp(bw, "ln1_g", f32(sd[a + "layer_norm.weight"])); setp(bw, "ln1_b", f32(sd[a + "layer_norm.bias"]))
            setp(bw, "w_c1", w_c1); setp(bw, "b_c1", f32(sd[f + "convs.0.conv.bias"]))
            setp(bw, "w_c2", w_c2); setp(bw, "b_c2", f32(sd[f + "convs.2.conv.bias"]))
            setp(bw, "ln2_g", f32(sd[f + "layer_norm.weight"])); setp(bw, "ln2_b", f32(sd[f + "layer_norm.bias"]))
            setp(bw, "w_qkv_bf", packed(w_qkv, 128, 64)); setp(bw, "w_out_bf", packed(w_out, 128, 64))
            setp(bw, "w_c1_bf", packed(w_c1, 256, 32)); setp(bw, "w_c2_bf", packed(w_c2, 128, 64))
            setp(bw, "w_ffn1_bf", packed(w_c1, 128, 64)); setp(bw, "w_ffn2_bf", packed(w_c2, 128, 64))

        p = "accent_encoder."
        for n in ("energy", "pitch"):
            setp(W, f"ae_{n}_w", f32(sd[p + f"{n}_embedding.conv.weight"].reshape(_lib.D, 3)))
            setp(W, f"ae_{n}_b", f32(sd[p + f"{n}_embedding.conv.bias"]))
        cfgs = [(256, 32, _lib.MEL_PAD), (256, 32, None), (128, 64, None)]
        for i, (nt, ks, kpad) in enumerate(cfgs):
            w = conv_w(p + f"convs.{4 * i}.conv.weight")
            setp(W, f"ae_c{i}_w", w); setp(W, f"ae_c{i}_b", f32(sd[p + f"convs.{4 * i}.conv.bias"]))
            setp(W, f"ae_ln{i}_g", f32(sd[p + f"convs.{4 * i + 2}.weight"]))
            setp(W, f"ae_ln{i}_b", f32(sd[p + f"convs.{4 * i + 2}.bias"]))
            setp(W, f"ae_c{i}_w_bf", packed(w, nt, ks, kpad))
        hp = self._hp
        W.n_ae_blocks = int(accent_encoder_params(hp)["nb_blocks"])
        W.n_pe_blocks = int(hp.phoneme_encoder["nb_blocks"])
        W.n_fd_blocks = int(hp.frame_decoder["nb_blocks"])
        for b in range(W.n_ae_blocks):
            fill_block(W.ae_blocks[b], f"accent_encoder.blocks.{b}")
        for b in range(W.n_pe_blocks):
            fill_block(W.pe_blocks[b], f"phoneme_encoder.blocks.{b}")
        for b in range(W.n_fd_blocks):
            fill_block(W.fd_blocks[b], f"frame_decoder.blocks.{b}")
        setp(W, "spk_w", f32(sd["spk_projection.linear_layer.weight"]))
        setp(W, "spk_b", f32(sd["spk_projection.linear_layer.bias"]))
        for i, k in enumerate((1, 3, 5)):
            setp(W, f"cls_w{i + 1}", f32(sd[f"speaker_classifier.classifier.{k}.linear_layer.weight"]))
            setp(W, f"cls_b{i + 1}", f32(sd[f"speaker_classifier.classifier.{k}.linear_layer.bias"]))
        setp(W, "film_g_w", f32(sd["style_adapter.gammas_predictor.linear_layer.weight"]))
        setp(W, "film_g_b", f32(sd["style_adapter.gammas_predictor.linear_layer.bias"]))
        setp(W, "film_b_w", f32(sd["style_adapter.betas_predictor.linear_layer.weight"]))
        setp(W, "film_b_b", f32(sd["style_adapter.betas_predictor.linear_layer.bias"]))
        setp(W, "post_mult", f32(sd["style_adapter.post_multipliers"]) if "style_adapter.post_multipliers" in sd else None)
        setp(W, "symbols_embedding", f32(sd["phoneme_encoder.symbols_embedding.weight"]))
        for n, short in (("duration", "dur"), ("energy", "energy"), ("pitch", "pitch")):
            setp(W, f"gu_{short}_w", f32(sd[f"gaussian_upsampling.{n}_projection.conv.weight"].reshape(_lib.D, 3)))
            setp(W, f"gu_{short}_b", f32(sd[f"gaussian_upsampling.{n}_projection.conv.bias"]))
        setp(W, "gu_range_w", f32(sd["gaussian_upsampling.projection.0.linear_layer.weight"].reshape(_lib.D)))
        setp(W, "gu_range_b", f32(sd["gaussian_upsampling.projection.0.linear_layer.bias"]))
        mel_w = lin_w("frame_decoder.projection.linear_layer.weight")
        setp(W, "mel_w", mel_w); setp(W, "mel_b", f32(sd["frame_decoder.projection.linear_layer.bias"]))
        setp(W, "mel_w_bf", packed(mel_w, 128, 64))
        pos = f32(_pos_table(_lib.D))       # built on the host with the reference's ops, uploaded once
        setp(W, "pos_table", pos)
        W.n_symbols = int(sd["phoneme_encoder.symbols_embedding.weight"].shape[0])
        W.n_speakers = int(self.n_speakers)
        W.pos_rows = int(pos.shape[0])
        self._prepared = (key, W, keep)
        return W

    def _workspace(self, B, L_max, T_max, slot=0):
        """One grow-only scratch buffer per slot (= per concurrent ``dx_forward`` of a call), reused for every shape
        that needs no more bytes.  ``torch.empty``: the kernels zero what they rely on (``zero_plan_halos``), and an
        allocation never enqueues a memset that could land under a running group.  A buffer that has to grow is
        replaced; the old block is returned to torch's caching allocator, which keeps it away from other streams
        until the kernels already enqueued on this stream have passed it (stream-ordered reuse)."""
        mode = _MODES[self.compute_mode]
        n = _lib.load().dx_forward_workspace_bytes(B, L_max, T_max, mode)
        if n < 0:
            _lib.check(1, "dx_forward_workspace_bytes")
        ws = self._workspaces.get(slot)
        if ws is None or ws.numel() < n or ws.device != self.device:
            ws = torch.empty(n, dtype=torch.uint8, device=self.device)
            self._workspaces[slot] = ws
        return ws

    # ------------------------------------------------------------------------------------------
    # reference API
    # ------------------------------------------------------------------------------------------
    def parse_batch(self, device, batch):
        """``DaftExprt.parse_batch`` (reference model.py:858-887): 14-tuple → (12 inputs, 6 targets)."""
        if len(batch) != 14:
            raise ValueError(
                f"Batch must have 14 elements (including speaker embeddings). Got {len(batch)}. "
                "Run training.py pre_process to compute ECAPA embeddings and ensure .spk_emb.npy files exist.")
        (symbols, durations_float, durations_int, symbols_energy, symbols_pitch, input_lengths,
         frames_energy, frames_pitch, mel_specs, output_lengths, speaker_ids, feature_dirs, feature_files, spk_embs) = batch
        to = lambda t, dt: t.to(device, non_blocking=True).to(dt)   # noqa: E731
        spk_embs = to(spk_embs, torch.float32)
        symbols = to(symbols, torch.long)
        durations_float = to(durations_float, torch.float32)
        durations_int = to(durations_int, torch.long)
        symbols_energy = to(symbols_energy, torch.float32)
        symbols_pitch = to(symbols_pitch, torch.float32)
        input_lengths = to(input_lengths, torch.long)
        frames_energy = to(frames_energy, torch.float32)
        frames_pitch = to(frames_pitch, torch.float32)
        mel_specs = to(mel_specs, torch.float32)
        output_lengths = to(output_lengths, torch.long)
        speaker_ids = to(speaker_ids, torch.long)
        inputs = (symbols, durations_float, durations_int, symbols_energy, symbols_pitch, input_lengths,
                  frames_energy, frames_pitch, mel_specs, output_lengths, speaker_ids, spk_embs)
        targets = (durations_float, symbols_energy, symbols_pitch, mel_specs, output_lengths, speaker_ids)
        return inputs, targets

    def _dev(self, t, dtype):
        """Tensor on the model's device, right dtype, contiguous (no copy when already so)."""
        if t is None:
            return None
        return t.to(device=self.device, dtype=dtype).contiguous()

    @staticmethod
    def _expect(name, t, shape):
        """Shape contract of one argument (``None`` passes).  The reference fails with a torch shape error when a
        tensor does not fit (or broadcasts a size-1 batch dimension); raw device pointers would read out of bounds."""
        if t is not None and tuple(t.shape) != tuple(shape):
            raise ValueError(f"{name}: expected shape {tuple(shape)}, got {tuple(t.shape)}")

    def _batch_rows(self, name, t, B, width):
        """(B, width) embedding; a (1, width) tensor is broadcast over the batch like the reference's
        ``accent_emb + spk_emb`` does (model.py:919)."""
        if t is None:
            return None
        if t.dim() == 2 and t.shape[0] == 1 and B > 1 and t.shape[1] == width:
            t = t.expand(B, width)
        self._expect(name, t, (B, width))
        return t.contiguous()

    def _run(self, B, L_max, T_max, *, symbols, durations_float, durations_int, symbols_energy, symbols_pitch,
             input_lengths, frames_energy=None, frames_pitch=None, mel_specs=None, output_lengths=None,
             spk_embs=None, accent_emb_in=None, spk_emb_in=None, want_enc_out=False, packed_frames=None):
        """The whole pass = ``dx_forward``.  Returns dict of output tensors.

        With ``self.streams > 1`` the batch is cut into that many contiguous groups of utterances which run as
        independent ``dx_forward`` calls on separate CUDA streams (joined before returning).  Utterances are
        independent given the padded widths ``L_max`` / ``T_max`` (they decide where the reference sees true
        zero padding) and every group is given the full batch's widths, so the results are
        bit-identical to the single call; the concurrency fills the tails of one group's kernels with the
        other group's work."""
        W = self._prepare()
        lib = _lib.load()
        dev = self.device
        for name, t in (("symbols", symbols), ("durations_float", durations_float), ("durations_int", durations_int),
                        ("symbols_energy", symbols_energy), ("symbols_pitch", symbols_pitch)):
            self._expect(name, t, (B, L_max))
        self._expect("input_lengths", input_lengths, (B,))
        self._expect("output_lengths", output_lengths, (B,))
        spk_embs = self._batch_rows("spk_embs", spk_embs, B, _lib.SPK)
        accent_emb_in = self._batch_rows("external_accent_emb", accent_emb_in, B, _lib.D)
        spk_emb_in = self._batch_rows("external_spk_emb", spk_emb_in, B, _lib.D)
        io_flags = 0
        if packed_frames is None:
            self._expect("frames_energy", frames_energy, (B, T_max))
            self._expect("frames_pitch", frames_pitch, (B, T_max))
            self._expect("mel_specs", mel_specs, (B, _lib.MEL, T_max))
            mel_shape = (B, _lib.MEL, T_max)
        else:
            # packed frame-side I/O (include/daft_b200.h DX_IO_PACKED_*): only the valid frames exist
            total = int(packed_frames)
            io_flags = _lib.IO_PACKED_MEL_OUT
            if mel_specs is not None:
                io_flags |= _lib.IO_PACKED_FRAMES
                self._expect("frames_energy (packed)", frames_energy, (total,))
                self._expect("frames_pitch (packed)", frames_pitch, (total,))
                self._expect("mel_specs (packed)", mel_specs, (total * _lib.MEL,))
            mel_shape = (total * _lib.MEL,)
        out = {
            "accent_emb": torch.empty(B, _lib.D, device=dev),
            "speaker_preds": torch.empty(B, self.n_speakers, device=dev),
            "film": torch.empty(B, 8, 2 * _lib.D, device=dev),
            "mel": torch.empty(mel_shape, device=dev),
            "weights": torch.empty(B, L_max, T_max, device=dev),
            "output_lengths": torch.empty(B, dtype=torch.long, device=dev),
            "enc_out": torch.empty(B, L_max, _lib.D, device=dev) if want_enc_out else None,
        }
        tensors = (("symbols", symbols), ("durations_float", durations_float), ("durations_int", durations_int),
                   ("symbols_energy", symbols_energy), ("symbols_pitch", symbols_pitch),
                   ("input_lengths", input_lengths), ("frames_energy", frames_energy),
                   ("frames_pitch", frames_pitch), ("mel_specs", mel_specs), ("output_lengths", output_lengths),
                   ("spk_embs", spk_embs), ("accent_emb_in", accent_emb_in), ("spk_emb_in", spk_emb_in),
                   ("accent_emb", out["accent_emb"]), ("speaker_preds", out["speaker_preds"]),
                   ("film", out["film"]), ("enc_out", out["enc_out"]), ("mel_out", out["mel"]),
                   ("weights_out", out["weights"]), ("output_lengths_out", out["output_lengths"]))
        n = max(1, min(int(self.streams), B // 8)) if (self.compute_mode == "bf16" and packed_frames is None) else 1
        bounds = [(B * i) // n for i in range(n + 1)]
        cur = torch.cuda.current_stream(dev)
        while len(self._side_streams) < n - 1:
            self._side_streams.append(torch.cuda.Stream(dev))
        # every group's scratch exists BEFORE the fork event is recorded: nothing the allocator enqueues on the
        # current stream can land under a group that is already running on a side stream
        wss = [self._workspace(bounds[i + 1] - bounds[i], L_max, T_max, slot=i) for i in range(n)]
        fork = cur.record_event() if n > 1 else None
        with torch.cuda.device(dev):
            for i in range(n):
                b0, b1 = bounds[i], bounds[i + 1]
                stream = cur if i == 0 else self._side_streams[i - 1]
                ws = wss[i]
                a = _lib.DxForwardArgs()
                a.B, a.L_max, a.T_max, a.mode = b1 - b0, L_max, T_max, _MODES[self.compute_mode]
                a.io_flags = io_flags
                for name, t in tensors:
                    setattr(a, name, None if t is None else (t.data_ptr() if n == 1 else t[b0:b1].data_ptr()))
                a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
                if i > 0:
                    stream.wait_event(fork)
                    ws.record_stream(stream)
                _lib.check(lib.dx_forward(C.byref(W), C.byref(a), C.c_void_p(stream.cuda_stream)), "dx_forward")
        for i in range(1, n):
            cur.wait_event(self._side_streams[i - 1].record_event())
        return out

    @torch.no_grad()
    def forward(self, inputs, external_accent_emb=None, external_spk_emb=None, _want_enc_out=False, packed_io=None):
        """``DaftExprt.forward`` (reference model.py:889-948); same 5-tuple.

        ``packed_io=(T_max, total_frames)`` (an extension for bulk generation, ``fine_tune.py:62-129``): the frame-side
        inputs ``frames_energy`` / ``frames_pitch`` / ``mel_specs`` are the PACKED tensors of ``layouts.pack_frames``
        (float[ΣT] / float[80·ΣT], per-utterance ``(80, T_b)`` blocks) and the returned mel is packed the same way
        (``layouts.unpack_mel`` / ``layouts.mel_views`` give the per-utterance crops ``fine_tune.py:79`` takes), so
        only valid frames cross PCIe.  Both numbers come from the host-side lengths the collate already has; no sync."""
        if len(inputs) != 12:
            raise ValueError(f"inputs must have 12 elements (including spk_embs). Got {len(inputs)}.")
        (symbols, durations_float, durations_int, symbols_energy, symbols_pitch, input_lengths,
         frames_energy, frames_pitch, mel_specs, output_lengths, speaker_ids, spk_embs) = inputs
        if external_spk_emb is None and spk_embs is None:
            raise ValueError("Speaker embeddings (spk_embs) required. Precompute ECAPA and provide .spk_emb.npy in data.")
        B, L_max = symbols.shape
        # T_max: the reference sizes frames by max(cumsum(durations_int)) (model.py:497), which the data
        # contract ties to the mel width (data_loader.py:165-168); the tensor shape avoids a host sync.
        if packed_io is not None:
            T_max, total_frames = int(packed_io[0]), int(packed_io[1])
        else:
            T_max = mel_specs.shape[2] if mel_specs is not None else int(output_lengths.max())
            total_frames = None
        fl, lg = torch.float32, torch.long
        out = self._run(
            B, L_max, T_max,
            symbols=self._dev(symbols, lg), durations_float=self._dev(durations_float, fl),
            durations_int=self._dev(durations_int, lg), symbols_energy=self._dev(symbols_energy, fl),
            symbols_pitch=self._dev(symbols_pitch, fl), input_lengths=self._dev(input_lengths, lg),
            frames_energy=None if external_accent_emb is not None else self._dev(frames_energy, fl),
            frames_pitch=None if external_accent_emb is not None else self._dev(frames_pitch, fl),
            mel_specs=None if external_accent_emb is not None else self._dev(mel_specs, fl),
            output_lengths=self._dev(output_lengths, lg),
            spk_embs=None if external_spk_emb is not None else self._dev(spk_embs, fl),
            accent_emb_in=self._dev(external_accent_emb, fl), spk_emb_in=self._dev(external_spk_emb, fl),
            want_enc_out=_want_enc_out, packed_frames=total_frames)
        n_pe, n_fd = int(self._hp.phoneme_encoder["nb_blocks"]), int(self._hp.frame_decoder["nb_blocks"])
        film_dec = out["film"][:, n_pe:n_pe + n_fd, :]
        film_params = [self.style_adapter.post_multipliers, None, None, film_dec]
        encoder_preds = [durations_float, symbols_energy, symbols_pitch, input_lengths]
        decoder_preds = [out["mel"], output_lengths]
        if _want_enc_out:
            self._last_debug = out
        return out["speaker_preds"], film_params, encoder_preds, decoder_preds, out["weights"]

    @torch.no_grad()
    def _accent_encoder(self, frames_energy, frames_pitch, mel_specs, output_lengths):
        """``AccentEncoder.forward`` (reference model.py:679-716) → (B, 128)."""
        W = self._prepare()
        lib = _lib.load()
        if mel_specs.dim() != 3 or mel_specs.shape[1] != _lib.MEL:
            raise ValueError(f"mel_specs: expected (B, {_lib.MEL}, T), got {tuple(mel_specs.shape)}")
        B, _, T_max = mel_specs.shape
        self._expect("frames_energy", frames_energy, (B, T_max))
        self._expect("frames_pitch", frames_pitch, (B, T_max))
        self._expect("output_lengths", output_lengths, (B,))
        ws = self._workspace(B, 1, T_max)
        fl = torch.float32
        fe, fp, mel = self._dev(frames_energy, fl), self._dev(frames_pitch, fl), self._dev(mel_specs, fl)
        ol = self._dev(output_lengths, torch.long)
        acc = torch.empty(B, _lib.D, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(lib.dx_accent_encoder(C.byref(W), _MODES[self.compute_mode], B, T_max, _lib.ptr(fe), _lib.ptr(fp),
                                             _lib.ptr(mel), _lib.ptr(ol), _lib.ptr(acc), _lib.ptr(ws), ws.numel(),
                                             _lib.stream_ptr(self.device)), "dx_accent_encoder")
        return acc

    # ---- inference-side prosody plumbing (reference model.py:950-1024), on the device ----------------
    def get_int_durations(self, duration_preds, hparams):
        """``DaftExprt.get_int_durations`` (reference model.py:950-973) + ``duration_to_integer``
        (extract_features.py:69-125).  Thresholds ``duration_preds`` in place and returns
        ``(duration_preds, durations_int)``.  One thread per utterance, Python-order double arithmetic."""
        lib = _lib.load()
        if duration_preds.device != self.device or duration_preds.dtype != torch.float32 or not duration_preds.is_contiguous():
            raise ValueError("duration_preds must be a contiguous fp32 tensor on the model's device")
        B, L = duration_preds.shape
        dint = torch.empty(B, L, dtype=torch.long, device=self.device)
        status = torch.empty(B, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(lib.dx_int_durations(_lib.ptr(duration_preds), B, L, int(hparams.sampling_rate),
                                            int(hparams.filter_length), int(hparams.hop_length),
                                            1 if hparams.centered else 0, _lib.ptr(dint), _lib.ptr(status),
                                            _lib.stream_ptr(self.device)), "dx_int_durations")
        self._last_status = status
        return duration_preds, dint

    def _pitch_stats(self, hparams, speaker_ids):
        """Per-utterance (mean, std) of log-pitch from ``hparams.stats['spk k']`` (reference model.py:983-985),
        gathered on the device from a small table so no ``.item()`` per row is needed."""
        if self._stats_table is None or self._stats_table[0] is not hparams.stats:
            ids = sorted(int(k.split()[1]) for k in hparams.stats if k.startswith("spk "))
            n = (max(ids) + 1) if ids else 1
            tab = torch.full((n, 2), float("nan"))
            for i in ids:
                st = hparams.stats[f"spk {i}"]["pitch"]
                tab[i, 0], tab[i, 1] = st["mean"], st["std"]
            self._stats_table = (hparams.stats, tab.to(self.device))
        tab = self._stats_table[1]
        ids = speaker_ids.to(self.device).long()
        # an id without an entry in hparams.stats is a KeyError in the reference (model.py:984).  The gather clamps a COPY of
        # the ids (the caller's tensor is never modified); the caller checks the ids themselves on the host - for free when
        # they are a CPU tensor, else folded into the host read it does anyway - with _raise_unknown_speakers.
        sel = tab.index_select(0, ids.clamp(0, tab.shape[0] - 1))
        return sel[:, 0].contiguous(), sel[:, 1].contiguous(), ids

    @staticmethod
    def _raise_unknown_speakers(hparams, ids_host):
        known = hparams.stats
        for i in ids_host.tolist():
            if f"spk {int(i)}" not in known:
                raise KeyError(f"spk {int(i)}")

    def _transform(self, energy, pitch, energy_factors, pitch_factors, durations_int, mean, std, transform):
        lib = _lib.load()
        B, L = pitch.shape
        for name, t in (("energy", energy), ("energy_factors", energy_factors), ("pitch_factors", pitch_factors),
                        ("durations_int", durations_int)):
            self._expect(name, t, (B, L))
        with torch.cuda.device(self.device):
            _lib.check(lib.dx_prosody_transform(_lib.ptr(energy), _lib.ptr(pitch), _lib.ptr(energy_factors),
                                                _lib.ptr(pitch_factors), _lib.ptr(durations_int), _lib.ptr(mean),
                                                _lib.ptr(std), B, L, transform, _lib.stream_ptr(self.device)),
                       "dx_prosody_transform")

    @torch.no_grad()
    def inference(self, inputs, pitch_transform, hparams, external_prosody=None, external_embeddings=None,
                  external_accent_emb=None):
        """``DaftExprt.inference`` (reference model.py:1026-1114); same 3-tuple, same exceptions."""
        symbols, dur_factors, energy_factors, pitch_factors, input_lengths, speaker_ids = inputs
        if external_embeddings is None:
            raise ValueError("external_embeddings required for inference. Provide ECAPA speaker embedding.")
        if external_accent_emb is None:
            raise ValueError("external_accent_emb required for inference. Provide --accent_emb_audios_dir or use a "
                             "checkpoint with memorized_accent_emb (e.g. from adapt_accent).")
        if external_prosody is None:
            raise ValueError("external_prosody must be provided for inference as the internal predictor has been removed.")
        if pitch_transform not in ("add", "multiply"):
            raise NotImplementedError
        fl, lg = torch.float32, torch.long
        B, L_max = symbols.shape
        self._expect("input_lengths", input_lengths, (B,))
        self._expect("speaker_ids", speaker_ids, (B,))
        for name, t in (("dur_factors", dur_factors), ("energy_factors", energy_factors), ("pitch_factors", pitch_factors),
                        ("external_prosody['duration_preds']", external_prosody["duration_preds"]),
                        ("external_prosody['energy_preds']", external_prosody["energy_preds"]),
                        ("external_prosody['pitch_preds']", external_prosody["pitch_preds"])):
            self._expect(name, t, (B, L_max))
        symbols_d = self._dev(symbols, lg)
        in_len = self._dev(input_lengths, lg)
        # duration_preds * dur_factors is a fresh tensor in the reference too (model.py:1074)
        duration_preds = (self._dev(external_prosody["duration_preds"], fl) * self._dev(dur_factors, fl)).contiguous()
        duration_preds, durations_int = self.get_int_durations(duration_preds, hparams)
        energy_preds = self._dev(external_prosody["energy_preds"], fl).clone()       # fresh (model.py:1077)
        pitch_src = external_prosody["pitch_preds"]
        pitch_preds = self._dev(pitch_src, fl)
        if pitch_preds.data_ptr() != pitch_src.data_ptr():
            pitch_preds = pitch_preds.clone()
        ids_dev = None
        if pitch_transform == "add":
            if speaker_ids.device.type == "cpu":
                self._raise_unknown_speakers(hparams, speaker_ids)          # host tensor: checked before any launch
                mean, std, _ = self._pitch_stats(hparams, speaker_ids)
            else:
                mean, std, ids_dev = self._pitch_stats(hparams, speaker_ids)
        else:
            mean = std = torch.zeros(B, device=self.device)
        self._transform(energy_preds, pitch_preds, self._dev(energy_factors, fl), self._dev(pitch_factors, fl),
                        durations_int, mean, std, 0 if pitch_transform == "add" else 1)
        # the only host sync of the pass: frame count (the reference syncs on it too, model.py:497,1100)
        # and the per-utterance status of the integer conversion
        rows = [durations_int.sum(dim=1), self._last_status.long()] + ([ids_dev] if ids_dev is not None else [])
        host = torch.stack(rows).cpu()
        if ids_dev is not None:
            self._raise_unknown_speakers(hparams, host[2])
        if bool((host[1] != 0).any()):
            bad = [int(i) for i in torch.nonzero(host[1]).flatten()]
            raise RuntimeError(f"get_int_durations: the reference's duration_to_integer fails for utterances {bad} "
                               "(frame / symbol count mismatch, extract_features.py:94-123, model.py:969)")
        T_max = int(host[0].max())
        assert T_max >= 1, "max(output_lengths) != T (reference model.py:1100): every duration is zero"
        out = self._run(
            B, L_max, T_max, symbols=symbols_d, durations_float=duration_preds, durations_int=durations_int,
            symbols_energy=energy_preds, symbols_pitch=pitch_preds, input_lengths=in_len,
            output_lengths=None, spk_embs=self._dev(external_embeddings, fl),
            accent_emb_in=self._dev(external_accent_emb, fl))
        if pitch_preds.data_ptr() != pitch_src.data_ptr() and pitch_src.dtype == fl:
            pitch_src.copy_(pitch_preds)          # the reference mutates the caller's tensor (H8)
        encoder_preds = [duration_preds, durations_int, energy_preds, pitch_preds, input_lengths]
        decoder_preds = [out["mel"], out["output_lengths"]]
        return encoder_preds, decoder_preds, out["weights"]

    def pitch_shift(self, pitch_preds, pitch_factors, hparams, speaker_ids):
        """``DaftExprt.pitch_shift`` (reference model.py:975-994), in place on a device tensor."""
        self._raise_unknown_speakers(hparams, speaker_ids.cpu())
        mean, std, _ = self._pitch_stats(hparams, speaker_ids)
        B, L = pitch_preds.shape
        ones = torch.ones(B, L, device=self.device)
        dint = torch.ones(B, L, dtype=torch.long, device=self.device)
        scratch = torch.zeros(B, L, device=self.device)
        self._transform(scratch, pitch_preds, ones, self._dev(pitch_factors, torch.float32), dint, mean, std, 0)
        return pitch_preds

    def pitch_multiply(self, pitch_preds, pitch_factors):
        """``DaftExprt.pitch_multiply`` (reference model.py:996-1024), in place on a device tensor."""
        B, L = pitch_preds.shape
        ones = torch.ones(B, L, device=self.device)
        dint = torch.ones(B, L, dtype=torch.long, device=self.device)
        scratch = torch.zeros(B, L, device=self.device)
        z = torch.zeros(B, device=self.device)
        self._transform(scratch, pitch_preds, ones, self._dev(pitch_factors, torch.float32), dint, z, z, 1)
        return pitch_preds
