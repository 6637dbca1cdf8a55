"""HiFi-GAN generator on B200, batched — mirror of the reference's ``daft_exprt/vocoder/hifigan.py``.

Same names and call contract as the reference:

* ``HiFiGANGenerator(config)`` (hifigan.py:125-184): ``load_state_dict`` takes the reference generator's checkpoint in
  either form (``weight_g`` / ``weight_v`` as saved, or plain ``weight`` after ``remove_weight_norm``);
  ``forward(x)`` maps a mel batch (B, 80, T) to (B, 1, T · 256).  Extra, ours: ``forward(x, lengths)`` vocodes every
  utterance as if alone on its cropped mel (what the reference achieves by calling ``infer`` once per utterance).
* ``HiFiGanVocoder(checkpoint_path, config, device)`` with ``infer(mel_spec)`` (hifigan.py:188-226): same input
  forms, same ``ValueError`` for a batch, same clipped numpy result; ``infer_batch`` is the batched entry point.
* ``load_hifigan_vocoder`` (hifigan.py:229-232).

The arithmetic runs in ``libdaft_b200.so`` (``dx_vocoder_forward``, csrc/dx_vocoder*.cu): every convolution is one
implicit-GEMM launch (tcgen05 in bf16 mode, SIMT in fp32 mode) with bias, residual, ResBlock averaging, leaky_relu
and the utterance-boundary zero padding fused into its epilogue.  There is no CPU fallback.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib

LRELU_SLOPE = 0.1
DEFAULT_CONFIG = {               # hifigan.py:28-41
    'sampling_rate': 22050,
    'upsample_rates': [8, 8, 2, 2],
    'upsample_kernel_sizes': [16, 16, 4, 4],
    'upsample_initial_channel': 512,
    'resblock': '1',
    'resblock_kernel_sizes': [3, 7, 11],
    'resblock_dilation_sizes': [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    'model_in_dim': 80,
}
_MODES = {"fp32": _lib.MODE_F32, "f32": _lib.MODE_F32, "bf16": _lib.MODE_BF16}


def _conv_specs(config):
    """[(prefix, kind, Cin, Cout, kernel, stride-or-dilation)] in the reference's registration order
    (hifigan.py:131-152)."""
    c0 = config['upsample_initial_channel']
    specs = [("conv_pre", "conv", config['model_in_dim'], c0, 7, 1)]
    for i, (u, k) in enumerate(zip(config['upsample_rates'], config['upsample_kernel_sizes'])):
        specs.append((f"ups.{i}", "convT", c0 // 2 ** i, c0 // 2 ** (i + 1), k, u))
    nk = len(config['resblock_kernel_sizes'])
    ch = c0
    for i in range(len(config['upsample_rates'])):
        ch = c0 // 2 ** (i + 1)
        for j, (k, dils) in enumerate(zip(config['resblock_kernel_sizes'], config['resblock_dilation_sizes'])):
            for l, d in enumerate(dils):
                specs.append((f"resblocks.{i * nk + j}.convs1.{l}", "conv", ch, ch, k, d))
            for l in range(len(dils)):
                specs.append((f"resblocks.{i * nk + j}.convs2.{l}", "conv", ch, ch, k, 1))
    specs.append(("conv_post", "conv", ch, 1, 7, 1))
    return specs


def synthetic_state_dict(config=None, seed=4321, weight_norm=True, gain=1.1):
    """Seeded random generator checkpoint (there is no network for the reference's default download, hifigan.py:43-68)
    in the reference's key layout, ``weight_g`` / ``weight_v`` form by default.  Layers have O(1) signal gain: the
    reference's own init N(0, 0.01²) (hifigan.py:73-74) would make every activation vanish.  Same sequence of draws as
    ``oracle/hifigan_oracle.make_state_dict`` (checked by tests/test_vocoder_oracle.py)."""
    import math
    config = DEFAULT_CONFIG if config is None else config
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, kind, cin, cout, k, _ in _conv_specs(config):
        std = gain / math.sqrt(cin * (k if kind == "conv" else 2))
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


def flops_per_frame(config=None):
    """Algorithmic FLOPs (2 · MAC) of the generator per mel frame: every Conv1d / ConvTranspose1d at its rate."""
    config = DEFAULT_CONFIG if config is None else config
    c0, rate, total = config['upsample_initial_channel'], 1, 2 * config['model_in_dim'] * config['upsample_initial_channel'] * 7
    for i, (u, k) in enumerate(zip(config['upsample_rates'], config['upsample_kernel_sizes'])):
        cin, cout = c0 // 2 ** i, c0 // 2 ** (i + 1)
        rate *= u
        total += 2 * cin * cout * (k // u) * rate                       # k / u taps reach one output sample
        for rk, dils in zip(config['resblock_kernel_sizes'], config['resblock_dilation_sizes']):
            total += 2 * cout * cout * rk * 2 * len(dils) * rate
    return total + 2 * cout * 7 * rate


def _check_config(config):
    if str(config.get('resblock', '1')) != '1':
        raise NotImplementedError("only ResBlock1 generators (config['resblock'] == '1', the reference default) are built")
    ups, ks = config['upsample_rates'], config['upsample_kernel_sizes']
    if len(ups) > 4 or len(config['resblock_kernel_sizes']) > 3 or any(len(d) > 3 for d in config['resblock_dilation_sizes']):
        raise NotImplementedError("at most 4 upsampling stages, 3 ResBlocks per stage and 3 layers per ResBlock")
    if len({len(d) for d in config['resblock_dilation_sizes']}) != 1:
        raise NotImplementedError("every ResBlock must have the same number of layers")
    for u, k in zip(ups, ks):
        if k != 2 * u or u < 2 or u & (u - 1):
            raise NotImplementedError(f"upsampler kernel {k} / stride {u}: only kernel == 2 * stride, stride a power of two")
    for k, dils in zip(config['resblock_kernel_sizes'], config['resblock_dilation_sizes']):
        if k % 2 != 1 or max(dils) * (k // 2) > _lib.VC_HALO:
            raise NotImplementedError(f"ResBlock kernel {k} with dilations {dils} exceeds the {_lib.VC_HALO}-row halo")
    if config['model_in_dim'] % 16 or (config['upsample_initial_channel'] // 2 ** len(ups)) % 32:
        raise NotImplementedError("mel channels must be a multiple of 16 and the last stage a multiple of 32 channels")


def polyphase_weight(w, u):
    """ConvTranspose1d weight (Cin, Cout, 2u) with stride u and padding u/2 (hifigan.py:135-142) → the 3-tap
    convolution at the INPUT rate it equals, GEMM layout [3][Cin][u * Cout]: output sample i·u + ph is column block ph
    of input-rate row i and only sees inputs i-1, i (ph < u/2) or i, i+1 (ph >= u/2):
    out[i·u + ph] = Σ_t x[i + t - 1] · g[t, :, ph, :]."""
    cin, cout, k = w.shape
    assert k == 2 * u
    g = torch.zeros(3, cin, u, cout, device=w.device, dtype=w.dtype)
    for ph in range(u):
        g[1, :, ph, :] = w[:, :, ph + u // 2]
        if ph < u // 2:
            g[0, :, ph, :] = w[:, :, ph + u // 2 + u]
        else:
            g[2, :, ph, :] = w[:, :, ph - u // 2]
    return g.reshape(3, cin, u * cout).contiguous()


class HiFiGANGenerator(torch.nn.Module):
    """Drop-in for the reference ``HiFiGANGenerator`` (inference only)."""

    def __init__(self, config):
        super().__init__()
        _check_config(config)
        self.config = config
        self.num_kernels = len(config['resblock_kernel_sizes'])
        self.num_upsamples = len(config['upsample_rates'])
        self.hop = int(np.prod(config['upsample_rates']))
        self.compute_mode = "bf16"
        self.residual = "bf16"           # bf16 mode only: dtype of the residual stream x ("fp32": tighter, 8 % slower)
        # folded (weight-norm-free) parameters under the reference's names: <prefix>.weight / <prefix>.bias
        for name, kind, cin, cout, k, _ in _conv_specs(config):
            shape = (cout, cin, k) if kind == "conv" else (cin, cout, k)
            self.register_buffer(name.replace(".", "__") + "__weight", torch.zeros(shape))
            self.register_buffer(name.replace(".", "__") + "__bias", torch.zeros(cout))
        self._prepared = None
        self._ws = {}
        # a pass needs ~160 KB of workspace per frame (bf16 mode; 4 bytes more per element in fp32 mode): larger
        # batches are vocoded in several passes over groups of utterances (identical results, tests/test_gpu_vocoder.py)
        self.max_frames_per_pass = 65536

    # ---- checkpoint -------------------------------------------------------------------------------------------
    def state_dict(self, *a, **k):
        sd = super().state_dict(*a, **k)
        return type(sd)((key.replace("__", "."), v) for key, v in sd.items())

    def load_state_dict(self, state_dict, strict=True):
        """Accepts the reference generator's ``state_dict`` with weight norm (``weight_g`` / ``weight_v``,
        folded as ``remove_weight_norm`` does, hifigan.py:178-184) or without."""
        sd = {(k[7:] if k.startswith("module.") else k): v for k, v in state_dict.items()}
        missing, used = [], set()
        for name, _, _, _, _, _ in _conv_specs(self.config):
            wbuf = getattr(self, name.replace(".", "__") + "__weight")
            bbuf = getattr(self, name.replace(".", "__") + "__bias")
            if name + ".weight_v" in sd:
                v, g = sd[name + ".weight_v"].float(), sd[name + ".weight_g"].float()
                w = torch._weight_norm(v.cpu(), g.cpu(), 0)
                used.update((name + ".weight_v", name + ".weight_g"))
            elif name + ".weight" in sd:
                w = sd[name + ".weight"].float()
                used.add(name + ".weight")
            else:
                missing.append(name + ".weight")
                continue
            if tuple(w.shape) != tuple(wbuf.shape):
                raise RuntimeError(f"size mismatch for {name}.weight: {tuple(w.shape)} vs {tuple(wbuf.shape)}")
            wbuf.copy_(w)
            if name + ".bias" in sd:
                bbuf.copy_(sd[name + ".bias"].float())
                used.add(name + ".bias")
            else:
                missing.append(name + ".bias")
        unexpected = [k for k in sd if k not in used]
        if strict and (missing or unexpected):
            raise RuntimeError(f"Error(s) in loading state_dict for HiFiGANGenerator: missing {missing}, unexpected {unexpected}")
        self._prepared = None
        return torch.nn.modules.module._IncompatibleKeys(missing, unexpected)

    def remove_weight_norm(self):
        """No-op: weight norm is folded when the checkpoint is loaded."""

    def set_mode(self, mode):
        if mode not in _MODES:
            raise ValueError(f"mode {mode!r}: expected 'fp32' or 'bf16'")
        if mode != self.compute_mode:
            self.compute_mode = mode
            self._prepared = None
        return self

    def set_residual(self, dtype):
        """bf16 mode: keep the residual stream in "bf16" (default; rel. L1 ≈ 1.2e-2 against the reference on the golden
        cases) or "fp32" (≈ 1.06e-2, 8 % slower).  fp32 mode is fp32 throughout."""
        if dtype not in ("bf16", "fp32"):
            raise ValueError(f"residual {dtype!r}: expected 'bf16' or 'fp32'")
        self.residual = dtype
        return self

    def _apply(self, fn, *a, **k):
        self._prepared = None
        self._ws = {}
        return super()._apply(fn, *a, **k)

    # ---- device-side weight layouts ---------------------------------------------------------------------------
    def _prepare(self):
        dev = getattr(self, "conv_pre__weight").device
        if dev.type != "cuda":
            raise _lib.DxError("HiFiGANGenerator (B200) needs its parameters on a CUDA device; there is no CPU fallback")
        if self._prepared is not None:
            return self._prepared
        lib = _lib.load()
        bf = _MODES[self.compute_mode] == _lib.MODE_BF16
        keep = []
        W = _lib.DxVocoderWeights()

        def fill(dst, name, kind, stride, tensor_path=True):
            w = getattr(self, name.replace(".", "__") + "__weight").float()
            b = getattr(self, name.replace(".", "__") + "__bias").float().contiguous()
            if kind == "conv":                       # (Cout, Cin, k) → [k][Cin][Cout]
                g = w.permute(2, 1, 0).contiguous()
            else:
                g = polyphase_weight(w, stride)
            taps, K, N = g.shape
            keep.extend((g, b))
            dst.w, dst.bias = g.data_ptr(), b.data_ptr()
            dst.taps, dst.K, dst.N = taps, K, N
            if bf and tensor_path:
                nt, ks = C.c_int(), C.c_int()
                _lib.check(lib.dx_vocoder_conv_tiling(taps, dst.dil, K, N, C.byref(nt), C.byref(ks)), "dx_vocoder_conv_tiling")
                n = lib.dx_packed_weight_elems(taps, K, N, nt.value, ks.value)
                out = torch.empty(n, dtype=torch.bfloat16, device=dev)
                _lib.check(lib.dx_pack_weight_bf16(_lib.ptr(g), taps, K, N, nt.value, ks.value, _lib.ptr(out), _lib.stream_ptr()),
                           "dx_pack_weight_bf16")
                keep.append(out)
                dst.w_bf = out.data_ptr()

        cfg = self.config
        W.n_stages, W.n_kernels = self.num_upsamples, self.num_kernels
        W.n_layers = len(cfg['resblock_dilation_sizes'][0])
        W.mel_channels = cfg['model_in_dim']
        c0 = cfg['upsample_initial_channel']
        W.pre.dil = 1
        fill(W.pre, "conv_pre", "conv", 1)
        for i, u in enumerate(cfg['upsample_rates']):
            W.up[i] = u
            W.channels[i] = c0 // 2 ** (i + 1)
            W.ups[i].dil = 1
            fill(W.ups[i], f"ups.{i}", "convT", u)
            for j, dils in enumerate(cfg['resblock_dilation_sizes']):
                for l, d in enumerate(dils):
                    W.rb[i][j][l][0].dil = d
                    fill(W.rb[i][j][l][0], f"resblocks.{i * self.num_kernels + j}.convs1.{l}", "conv", 1)
                    W.rb[i][j][l][1].dil = 1
                    fill(W.rb[i][j][l][1], f"resblocks.{i * self.num_kernels + j}.convs2.{l}", "conv", 1)
        W.post.dil = 1
        fill(W.post, "conv_post", "conv", 1, tensor_path=False)     # [7][C][1]: the SIMT tail kernel
        self._prepared = (W, keep)
        return self._prepared

    def _workspace(self, W, B, frame_cap):
        key = (B, frame_cap, self.compute_mode)
        ws = self._ws.get(key)
        if ws is None:
            n = _lib.load().dx_vocoder_workspace_bytes(C.byref(W), B, frame_cap, _MODES[self.compute_mode])
            if n < 0:
                _lib.check(1, "dx_vocoder_workspace_bytes")
            if len(self._ws) > 4:
                self._ws.clear()
            ws = self._ws[key] = torch.empty(n, dtype=torch.uint8, device=getattr(self, "conv_pre__weight").device)
        return ws

    # ---- forward ----------------------------------------------------------------------------------------------
    def forward(self, x, lengths=None, total_frames=None):
        """x (B, 80, T) → (B, 1, T · hop).  ``lengths`` (frames per utterance; tensor, list or None = all T): every
        utterance is vocoded on its cropped mel, samples behind it are 0.  ``total_frames`` (optional host int, the
        sum of lengths) sizes the packed stream without a device sync when ``lengths`` lives on the GPU."""
        W, _ = self._prepare()
        lib = _lib.load()
        dev = getattr(self, "conv_pre__weight").device
        if x.dim() != 3 or x.size(1) != self.config['model_in_dim']:
            raise RuntimeError(f"expected mel of shape (B, {self.config['model_in_dim']}, T), got {tuple(x.shape)}")
        x = x.to(device=dev, dtype=torch.float32).contiguous()
        B, _, T = x.shape
        host = None
        if lengths is None:
            total = B * T
            host = torch.full((B,), T, dtype=torch.int64)
            len_dev = torch.full((B,), T, dtype=torch.int64, device=dev)
        else:
            if torch.is_tensor(lengths) and lengths.is_cuda:
                total = int(total_frames) if total_frames is not None else B * T
                len_dev = lengths.to(dtype=torch.int64).contiguous()
            else:
                host = torch.as_tensor(lengths, dtype=torch.int64)
                if host.numel() != B or int(host.max()) > T or int(host.min()) < 0:
                    raise ValueError("lengths must hold one value in [0, T] per utterance")
                total = int(host.sum())
                len_dev = host.to(dev)
        audio = torch.empty(B, 1, T * self.hop, device=dev)
        if total > self.max_frames_per_pass and B > 1 and host is not None:
            # groups of consecutive utterances with at most max_frames_per_pass frames each
            b0 = 0
            while b0 < B:
                b1, acc = b0 + 1, int(host[b0])
                while b1 < B and acc + int(host[b1]) <= self.max_frames_per_pass:
                    acc += int(host[b1])
                    b1 += 1
                self._run(W, x[b0:b1], len_dev[b0:b1], b1 - b0, T, acc, audio[b0:b1])
                b0 = b1
        else:
            self._run(W, x, len_dev, B, T, total, audio)
        return audio

    def _run(self, W, x, len_dev, B, T, total, audio):
        lib = _lib.load()
        frame_cap = lib.dx_vocoder_frame_cap(B, total)
        ws = self._workspace(W, B, frame_cap)
        a = _lib.DxVocoderArgs()
        a.B, a.T_max, a.mode, a.frame_cap = B, T, _MODES[self.compute_mode], frame_cap
        a.flags = _lib.VC_RESIDUAL_F32 if self.residual == "fp32" else 0
        a.mel, a.lengths, a.audio = x.data_ptr(), len_dev.data_ptr(), audio.data_ptr()
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        _lib.check(lib.dx_vocoder_forward(C.byref(W), C.byref(a), _lib.stream_ptr()), "dx_vocoder_forward")
        return audio


class HiFiGanVocoder:
    """Drop-in for the reference ``HiFiGanVocoder`` (hifigan.py:187-226).  ``checkpoint_path=None`` makes the reference
    download its default checkpoint (hifigan.py:43-68); there is no network here, so a path or ``state_dict`` is
    required."""

    def __init__(self, checkpoint_path=None, config=None, device='cuda', state_dict=None, mode="bf16", residual="bf16"):
        self.device = torch.device(device)
        self.config = DEFAULT_CONFIG if config is None else config
        self.checkpoint_path = checkpoint_path
        if state_dict is None:
            if checkpoint_path is None or not os.path.isfile(checkpoint_path):
                raise FileNotFoundError("HiFi-GAN checkpoint not found: pass checkpoint_path (a local generator "
                                        "checkpoint) or state_dict; the default checkpoint cannot be downloaded here")
            checkpoint = torch.load(checkpoint_path, map_location='cpu')
            state_dict = checkpoint.get('generator', checkpoint.get('state_dict', checkpoint))     # hifigan.py:202
        self.generator = HiFiGANGenerator(self.config)
        self.generator.load_state_dict(state_dict)
        self.generator.to(self.device).eval().set_mode(mode).set_residual(residual)

    def infer(self, mel_spec):
        if isinstance(mel_spec, np.ndarray):
            mel = torch.from_numpy(mel_spec).float()
        elif torch.is_tensor(mel_spec):
            mel = mel_spec.float()
        else:
            mel = torch.tensor(mel_spec, dtype=torch.float32)
        if mel.dim() == 2:
            mel = mel.unsqueeze(0)
        if mel.dim() == 3 and mel.size(0) != 1:
            raise ValueError('Mel spectrogram batch inference not supported in this helper.')
        with torch.no_grad():
            audio = self.generator(mel.to(self.device)).squeeze().cpu().numpy()
        return np.clip(audio, -1.0, 1.0)

    def infer_batch(self, mels, lengths):
        """mels (B, 80, T_max) (host or device), lengths (B,) → list of clipped numpy waveforms, one per utterance;
        one pass over the packed batch instead of B calls."""
        mels = torch.as_tensor(mels)
        host_len = [int(v) for v in (lengths.tolist() if torch.is_tensor(lengths) else lengths)]
        with torch.no_grad():
            audio = self.generator(mels.to(self.device), torch.tensor(host_len)).squeeze(1)
            audio = audio.clamp_(-1.0, 1.0).cpu().numpy()
        hop = self.generator.hop
        return [audio[b, : n * hop] for b, n in enumerate(host_len)]


def load_hifigan_vocoder(checkpoint_path=None, device=None):
    if device is None:
        device = 'cuda' if torch.cuda.is_available() else 'cpu'
    return HiFiGanVocoder(checkpoint_path=checkpoint_path, config=DEFAULT_CONFIG, device=device)
