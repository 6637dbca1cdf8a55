"""Training step of BASELINE.json config 5: the reference's ``train.py:264-280,421-445`` iteration on B200.

    outputs = model(inputs); loss, individual = criterion(outputs, targets, iteration)     # train.py:421-422
    loss.backward()                                                                         # :434  (DDP all-reduces here, :272)
    optimizer.step()                                                                        # :445  (Adam, :278-280)

``TrainEngine`` owns the parameters as ONE flat fp32 buffer in the engine layout of ``include/daft_b200.h``
(conv weights ``[tap][Cin][Cout]``, in/out projections transposed, everything else as torch stores it), a flat
gradient buffer of the same layout and the two Adam moment buffers.  One step is

    dx_train_step      forward with saved activations + DaftExprtLoss + hand-written backward of every module (CUDA)
    all_reduce(SUM)    of the flat gradient buffer over NCCL when torch.distributed is initialised (the path's only
                       collective; averaging = the 1 / world_size folded into the Adam kernel)
    dx_adam_step       torch.optim.Adam semantics on the flat buffers (CUDA)

``state_dict()`` / ``load_state_dict()`` speak the reference's checkpoint layout (184 tensors), so a checkpoint moves
between the reference, the inference mirror (``model.DaftExprt``) and this engine unchanged.  Dropout is the identity
(the parity configuration of SURVEY.md section 8d; the reference's training runs p = 0.1).  There is no CPU fallback.
"""
import ctypes as C
import math

import torch

from . import _lib
from .hparams import accent_encoder_params
from .model import _pos_table
from .weights import state_dict_spec, strip_module_prefix

LOSS_NAMES = ("loss", "speaker_loss", "speaker_ce_raw", "post_mult_loss", "mel_spec_l1_loss", "mel_spec_l2_loss",
              "energy_consistency_loss", "pitch_consistency_loss")


def _block_fields(prefix):
    a, f = prefix + ".attention.", prefix + ".feed_forward."
    return [(a + "multi_head_attention.in_proj_weight", "w_qkv", "lin_t"), (a + "multi_head_attention.in_proj_bias", "b_qkv", "same"),
            (a + "multi_head_attention.out_proj.weight", "w_out", "lin_t"), (a + "multi_head_attention.out_proj.bias", "b_out", "same"),
            (a + "layer_norm.weight", "ln1_g", "same"), (a + "layer_norm.bias", "ln1_b", "same"),
            (f + "convs.0.conv.weight", "w_c1", "conv"), (f + "convs.0.conv.bias", "b_c1", "same"),
            (f + "convs.2.conv.weight", "w_c2", "conv"), (f + "convs.2.conv.bias", "b_c2", "same"),
            (f + "layer_norm.weight", "ln2_g", "same"), (f + "layer_norm.bias", "ln2_b", "same")]


def field_map():
    """state_dict name -> (struct path, transform).  struct path: ('ae_c0_w',) or ('pe_blocks', 2, 'w_qkv')."""
    m = {}
    p = "accent_encoder."
    for n in ("energy", "pitch"):
        m[p + f"{n}_embedding.conv.weight"] = ((f"ae_{n}_w",), "same")
        m[p + f"{n}_embedding.conv.bias"] = ((f"ae_{n}_b",), "same")
    for i in range(3):
        m[p + f"convs.{4 * i}.conv.weight"] = ((f"ae_c{i}_w",), "conv")
        m[p + f"convs.{4 * i}.conv.bias"] = ((f"ae_c{i}_b",), "same")
        m[p + f"convs.{4 * i + 2}.weight"] = ((f"ae_ln{i}_g",), "same")
        m[p + f"convs.{4 * i + 2}.bias"] = ((f"ae_ln{i}_b",), "same")
    for arr, mod in (("ae_blocks", "accent_encoder"), ("pe_blocks", "phoneme_encoder"), ("fd_blocks", "frame_decoder")):
        for b in range(4):
            for name, field, tr in _block_fields(f"{mod}.blocks.{b}"):
                m[name] = ((arr, b, field), tr)
    m["spk_projection.linear_layer.weight"] = (("spk_w",), "same")
    m["spk_projection.linear_layer.bias"] = (("spk_b",), "same")
    for i, k in enumerate((1, 3, 5)):
        m[f"speaker_classifier.classifier.{k}.linear_layer.weight"] = ((f"cls_w{i + 1}",), "same")
        m[f"speaker_classifier.classifier.{k}.linear_layer.bias"] = ((f"cls_b{i + 1}",), "same")
    m["style_adapter.gammas_predictor.linear_layer.weight"] = (("film_g_w",), "same")
    m["style_adapter.gammas_predictor.linear_layer.bias"] = (("film_g_b",), "same")
    m["style_adapter.betas_predictor.linear_layer.weight"] = (("film_b_w",), "same")
    m["style_adapter.betas_predictor.linear_layer.bias"] = (("film_b_b",), "same")
    m["style_adapter.post_multipliers"] = (("post_mult",), "same")
    m["phoneme_encoder.symbols_embedding.weight"] = (("symbols_embedding",), "same")
    for n, short in (("duration", "dur"), ("energy", "energy"), ("pitch", "pitch")):
        m[f"gaussian_upsampling.{n}_projection.conv.weight"] = ((f"gu_{short}_w",), "same")
        m[f"gaussian_upsampling.{n}_projection.conv.bias"] = ((f"gu_{short}_b",), "same")
    m["gaussian_upsampling.projection.0.linear_layer.weight"] = (("gu_range_w",), "same")
    m["gaussian_upsampling.projection.0.linear_layer.bias"] = (("gu_range_b",), "same")
    m["frame_decoder.projection.linear_layer.weight"] = (("mel_w",), "lin_t")
    m["frame_decoder.projection.linear_layer.bias"] = (("mel_b",), "same")
    return m


def to_engine(t, transform):
    if transform == "conv":            # torch [Cout][Cin][k] -> [k][Cin][Cout]
        return t.permute(2, 1, 0).contiguous()
    if transform == "lin_t":           # torch [out][in] -> [in][out]
        return t.t().contiguous()
    return t.contiguous()


def from_engine(flat, shape, transform):
    """Flat engine-layout slice -> tensor of the torch ``shape``."""
    if transform == "conv":
        co, ci, k = shape
        return flat.view(k, ci, co).permute(2, 1, 0).contiguous()
    if transform == "lin_t":
        o, i = shape
        return flat.view(i, o).t().contiguous()
    return flat.view(shape).clone()


def update_learning_rate(hparams, iteration):
    """``train.py:148-160``: linear warm-up, then inverse square root decay."""
    warmup, lr0, lr1 = hparams.warmup_steps, hparams.initial_learning_rate, hparams.max_learning_rate
    if iteration < warmup:
        return (lr1 - lr0) / warmup * iteration + lr0
    return iteration ** -0.5 * lr1 / warmup ** -0.5


def adversarial_weight(hparams, iteration):
    """``DaftExprtLoss.update_adversarial_weight`` (loss.py:51-55)."""
    warmup = getattr(hparams, "warmup_steps", 10000)
    adv_max = getattr(hparams, "adv_max_weight", 1e-2)
    return min(adv_max, iteration * warmup ** -1.5 * adv_max / warmup ** -0.5)


class TrainEngine:
    def __init__(self, hparams, state_dict, device="cuda", gemm="fp32", train_mode=False, seed=0):
        """``train_mode``: ``model.train()`` (train.py:411) - the dropout layers of the three stacks and of the accent
        encoder's pre-net are active, with the probabilities of ``hparams.{accent_encoder, phoneme_encoder,
        frame_decoder}``; needs ``gemm="bf16"``.  False = ``model.eval()``: dropout is the identity (the gradient-parity
        configuration).  ``seed``: the masks of an iteration are a pure function of (seed, iteration, micro-batch index).

        ``gemm``: "fp32" - every contraction in fp32 SIMT (the 1e-3 gradient-parity mode); "bf16" - the large dense
        contractions (forward convs / linears, their dgrad and wgrad) on tcgen05 with bf16 copies of the fp32 operands
        and fp32 accumulation (``DX_TRAIN_TC_GEMM``); LayerNorm, attention, length regulator, loss, reductions and Adam
        stay fp32 either way."""
        from .model import DaftExprt
        DaftExprt._check_hparams(hparams)
        if gemm not in ("fp32", "bf16"):
            raise ValueError("gemm must be 'fp32' or 'bf16'")
        self.gemm = gemm
        self.hp = hparams
        # FFT blocks per stack: accent encoder, phoneme encoder, frame decoder (1..4 each, checked above)
        self.n_blocks = tuple(int(x["nb_blocks"]) for x in (accent_encoder_params(hparams), hparams.phoneme_encoder,
                                                              hparams.frame_decoder))
        self.train_mode = bool(train_mode)
        self.seed = int(seed)
        if self.train_mode and gemm != "bf16":
            raise ValueError("train_mode=True (dropout) is implemented on the gemm='bf16' path only")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.DxError("TrainEngine needs a CUDA device; there is no CPU fallback")
        self.lib = _lib.load()
        self.spec = [(n, tuple(s)) for n, s in state_dict_spec(hparams)]
        if getattr(hparams, "post_mult_weight", 0.0) == 0.0:
            self.spec = [(n, s) for n, s in self.spec if n != "style_adapter.post_multipliers"]
        self.fmap = field_map()
        self.offsets, off = {}, 0
        for name, shape in self.spec:
            n = 1
            for d in shape:
                n *= d
            self.offsets[name] = (off, n)
            off += (n + 63) // 64 * 64                     # 256-byte aligned slices (float4 loads in the kernels)
        self.numel = off
        dev = self.device
        self.params = torch.zeros(off, device=dev)
        self.grads = torch.zeros(off, device=dev)
        self.exp_avg = torch.zeros(off, device=dev)
        self.exp_avg_sq = torch.zeros(off, device=dev)
        self.pos = _pos_table(_lib.D).to(dev)
        self.W = self._struct(self.params)
        self.G = self._struct(self.grads)
        self.losses = torch.zeros(8, device=dev)
        self.step_count = 0
        self._ws = None
        self.load_state_dict(state_dict)

    def _struct(self, flat):
        W = _lib.DxModelWeights()
        base = flat.data_ptr()
        for name, _ in self.spec:
            path, _tr = self.fmap[name]
            ptr = base + 4 * self.offsets[name][0]
            if len(path) == 1:
                setattr(W, path[0], ptr)
            else:
                setattr(getattr(W, path[0])[path[1]], path[2], ptr)
        W.pos_table = self.pos.data_ptr()
        W.n_symbols = int(dict(self.spec)["phoneme_encoder.symbols_embedding.weight"][0])
        W.n_speakers = int(self.hp.n_speakers)
        W.pos_rows = int(self.pos.shape[0])
        W.n_ae_blocks, W.n_pe_blocks, W.n_fd_blocks = self.n_blocks
        return W

    # ---- checkpoint layout in / out ---------------------------------------------------------------------------------
    def load_state_dict(self, state_dict):
        sd = strip_module_prefix(state_dict)
        for name, shape in self.spec:
            t = sd[name].detach().to(device=self.device, dtype=torch.float32)
            if tuple(t.shape) != shape:
                raise ValueError(f"{name}: expected shape {shape}, got {tuple(t.shape)}")
            off, n = self.offsets[name]
            self.params[off:off + n].copy_(to_engine(t, self.fmap[name][1]).reshape(-1))

    def _export(self, flat):
        return {name: from_engine(flat[self.offsets[name][0]:self.offsets[name][0] + self.offsets[name][1]], shape, self.fmap[name][1])
                for name, shape in self.spec}

    def state_dict(self):
        return self._export(self.params)

    def grads_state_dict(self):
        """The gradients in the reference's parameter layout (what ``p.grad`` holds after ``loss.backward()``)."""
        return self._export(self.grads)

    # ---- one iteration --------------------------------------------------------------------------------------------------
    def _workspace(self, B, L, T):
        n = self.lib.dx_train_workspace_bytes(B, L, T)
        if n < 0:
            _lib.check(1, "dx_train_workspace_bytes")
        if self._ws is None or self._ws.numel() < n:
            self._ws = torch.empty(n, dtype=torch.uint8, device=self.device)
        return self._ws

    def dropout_probabilities(self):
        """(attn_dropout[3], conv_dropout[3]) by stack: accent encoder, phoneme encoder, frame decoder."""
        stacks = (accent_encoder_params(self.hp), self.hp.phoneme_encoder, self.hp.frame_decoder)
        return ([float(s.get("attn_dropout", 0.0)) for s in stacks], [float(s.get("conv_dropout", 0.0)) for s in stacks])

    def dropout_seed(self, iteration, micro_batch=0):
        """64-bit seed of one forward pass: splitmix64 of (seed, iteration, micro-batch, rank), so that consecutive
        iterations and the ranks of a data-parallel job draw unrelated masks."""
        rank = 0
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                rank = dist.get_rank()
        except Exception:        # noqa: BLE001
            rank = 0
        m = (1 << 64) - 1
        z = (self.seed * 0x9E3779B97F4A7C15 + int(iteration) * 0xD1B54A32D192ED03 + int(micro_batch) * 0x8CB92BA72F3D8DD7
             + rank * 0xABC98388FB8FAC03 + 0x2545F4914F6CDD1D) & m
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & m
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & m
        return z ^ (z >> 31)

    def dropout_masks(self, B, L_max, T_max, iteration=1, micro_batch=0):
        """The keep-masks the next ``forward_backward(..., iteration, micro_batch=...)`` applies, in the ORDER the reference's
        forward reaches its dropout calls (model.py:889-948): accent encoder pre-net x 3, then per stack (accent encoder,
        phoneme encoder, frame decoder) and block: attention weights (B * 2, n, n), attention output (B, n, 128),
        feed-forward output (B, n, 128).  Returns a list of (bool tensor on the device, p).  Test / debugging surface."""
        pa, pc = self.dropout_probabilities()
        seed = self.dropout_seed(iteration, micro_batch)
        out = []

        def site(kind, stack, block, p, n, Cdim):
            shape = (B * 2, n, n) if kind == 3 else (B, n, Cdim)
            m = torch.empty(shape, dtype=torch.uint8, device=self.device)
            with torch.cuda.device(self.device):
                _lib.check(self.lib.dx_dropout_mask(seed, kind, stack, block, float(p), B, n, Cdim, _lib.ptr(m),
                                                    _lib.stream_ptr(self.device)), "dx_dropout_mask")
            out.append((m.bool(), float(p)))

        for k, Cdim in enumerate((_lib.CCH, _lib.CCH, _lib.D)):
            site(2, 0, k, pc[0], T_max, Cdim)
        for stack, n in ((0, T_max), (1, L_max), (2, T_max)):
            for blk in range(self.n_blocks[stack]):
                site(3, stack, blk, pa[stack], n, _lib.D)
                site(0, stack, blk, pa[stack], n, _lib.D)
                site(1, stack, blk, pc[stack], n, _lib.D)
        return out

    def forward_backward(self, inputs, iteration=1, loss_scale=1.0, forward_only=False, want_mel=False, micro_batch=0):
        """``inputs``: the 12-tuple of ``DaftExprt.forward`` (targets are a subset of it, model.py:884-885) on the device.
        Fills ``self.grads`` (this rank's gradients) and returns the 8 loss terms as a device tensor (no sync).
        ``forward_only`` is the validation pass (train.py:155-209, ``model.eval()``): dropout is off there."""
        (symbols, dfl, dint, sen, spi, in_len, fen, fpi, mel, out_len, spk_ids, spk_embs) = inputs
        B, L = symbols.shape
        T = mel.shape[2]
        fl, lg = torch.float32, torch.long
        t = [symbols.to(self.device, lg), dfl.to(self.device, fl), dint.to(self.device, lg), sen.to(self.device, fl),
             spi.to(self.device, fl), in_len.to(self.device, lg), fen.to(self.device, fl), fpi.to(self.device, fl),
             mel.to(self.device, fl), out_len.to(self.device, lg), spk_ids.to(self.device, lg), spk_embs.to(self.device, fl)]
        t = [x.contiguous() for x in t]
        shapes = [(B, L)] * 5 + [(B,), (B, T), (B, T), (B, _lib.MEL, T), (B,), (B,), (B, _lib.SPK)]
        for x, s, n in zip(t, shapes, ("symbols", "durations_float", "durations_int", "symbols_energy", "symbols_pitch",
                                       "input_lengths", "frames_energy", "frames_pitch", "mel_specs", "output_lengths",
                                       "speaker_ids", "spk_embs")):
            if tuple(x.shape) != s:
                raise ValueError(f"{n}: expected shape {s}, got {tuple(x.shape)}")
        ws = self._workspace(B, L, T)
        a = _lib.DxTrainArgs()
        a.B, a.L_max, a.T_max, a.flags = B, L, T, ((_lib.TRAIN_FORWARD_ONLY if forward_only else 0) |
                                                (_lib.TRAIN_TC_GEMM if self.gemm == "bf16" else 0))
        for name, x in zip(("symbols", "durations_float", "durations_int", "symbols_energy", "symbols_pitch", "input_lengths",
                            "frames_energy", "frames_pitch", "mel_specs", "output_lengths", "speaker_ids", "spk_embs"), t):
            setattr(a, name, x.data_ptr())
        hp = self.hp
        a.mel_spec_weight = float(getattr(hp, "mel_spec_weight", 1.0))
        a.post_mult_weight = float(getattr(hp, "post_mult_weight", 1e-3))
        a.energy_consistency_weight = float(getattr(hp, "energy_consistency_weight", 0.0))
        a.speaker_weight = float(adversarial_weight(hp, iteration))
        a.lambda_reversal = float(hp.lambda_reversal)
        a.loss_scale = float(loss_scale)
        a.losses = self.losses.data_ptr()
        mel_out = torch.empty(B, _lib.MEL, T, device=self.device) if want_mel else None
        a.mel_out = None if mel_out is None else mel_out.data_ptr()
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        if self.train_mode and not forward_only:
            pa, pc = self.dropout_probabilities()
            for k in range(3):
                a.attn_dropout[k], a.conv_dropout[k] = pa[k], pc[k]
            a.dropout_seed = self.dropout_seed(iteration, micro_batch)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dx_train_step(C.byref(self.W), C.byref(self.G), C.byref(a), _lib.stream_ptr(self.device)),
                       "dx_train_step")
        self._keep = t
        return (self.losses, mel_out) if want_mel else self.losses

    def all_reduce_grads(self):
        """DDP's gradient exchange (train.py:269-272) as ONE all-reduce of the flat buffer; the average's 1 / world_size
        is applied inside the Adam kernel.  Returns the world size."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.grads, op=dist.ReduceOp.SUM)
            return dist.get_world_size()
        return 1

    def adam(self, lr, world=1):
        hp = self.hp
        self.step_count += 1
        b1, b2 = getattr(hp, "betas", (0.9, 0.98))
        with torch.cuda.device(self.device):
            _lib.check(self.lib.dx_adam_step(_lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(self.exp_avg),
                                             _lib.ptr(self.exp_avg_sq), self.numel, float(lr), float(b1), float(b2),
                                             float(getattr(hp, "epsilon", 1e-9)), float(getattr(hp, "weight_decay", 1e-6)),
                                             int(self.step_count), 1.0 / world, _lib.stream_ptr(self.device)), "dx_adam_step")

    def step(self, inputs, iteration=1, lr=None):
        """One training iteration with ``accumulation_steps = 1`` and no gradient clipping (the defaults of the benchmark
        configuration).  Returns the loss terms (device tensor, no sync).  ``iterate`` is the general form."""
        losses = self.forward_backward(inputs, iteration)
        world = self.all_reduce_grads()
        self.adam(update_learning_rate(self.hp, iteration) if lr is None else lr, world)
        return losses

    def iterate(self, micro_batches, iteration=1, lr=None):
        """The reference's full iteration (train.py:421-445) over ``hparams.accumulation_steps`` micro-batches:
        ``loss / accumulation_steps`` per micro-batch (:423), gradients accumulated over them (:434-436), ONE all-reduce
        of the flat buffer (DDP averages at the last backward, :269-272), ``clip_grad_norm_(parameters,
        hparams.grad_clip_thresh)`` (:441), ``optimizer.step()`` (:445).  Returns (summed loss terms as a device tensor,
        gradient norm before clipping as a device scalar).  A finite ``grad_clip_thresh`` costs one host
        synchronisation per iteration (the reference has the same one: it logs the norm)."""
        micro_batches = list(micro_batches)
        n = len(micro_batches)
        want = int(getattr(self.hp, "accumulation_steps", 1))
        if n != want:
            raise ValueError(f"expected hparams.accumulation_steps = {want} micro-batches, got {n}")
        total = None
        acc = None
        for k, mb in enumerate(micro_batches):
            losses = self.forward_backward(mb, iteration, loss_scale=1.0 / n, micro_batch=k)
            total = losses.clone() if total is None else total + losses
            if n > 1:
                if acc is None:
                    acc = self.grads.clone()
                else:
                    acc.add_(self.grads)
        if acc is not None:
            self.grads.copy_(acc)
        world = self.all_reduce_grads()
        norm = torch.linalg.vector_norm(self.grads) / world         # norm of the AVERAGED gradient, as DDP leaves it
        scale = 1.0 / world
        thresh = float(getattr(self.hp, "grad_clip_thresh", float("inf")))
        if math.isfinite(thresh):
            coef = thresh / (float(norm) + 1e-6)                        # torch.nn.utils.clip_grad_norm_
            if coef < 1.0:
                scale *= coef
        self.adam(update_learning_rate(self.hp, iteration) if lr is None else lr, 1.0 / scale)
        return total, norm

    def grad_norm(self):
        """What ``clip_grad_norm_(parameters, inf)`` reports (train.py:441); padding between slices is zero."""
        return torch.linalg.vector_norm(self.grads)


def loss_dict(losses):
    v = losses.detach().cpu().tolist()
    return dict(zip(LOSS_NAMES, v))


def flops_per_step(in_lens, out_lens):
    """Algorithmic FLOPs of forward + backward (backward = 2 x forward for every contraction)."""
    D, CC, KT, M, E = 128, 1024, 3, 80, 192
    tot = 0.0
    for l, t in zip(in_lens, out_lens):
        blk = lambda n: 8 * n * D * D + 4 * n * n * D + 4 * n * KT * D * CC     # noqa: E731
        acc = 2 * t * KT * (M * CC + CC * CC + CC * D) + 4 * blk(t)
        tot += acc + 4 * blk(l) + 4 * blk(t) + 2 * t * D * M + 2 * l * t * D
    return 3.0 * tot


__all__ = ["TrainEngine", "update_learning_rate", "adversarial_weight", "loss_dict", "LOSS_NAMES", "field_map", "math"]
