"""GPU: unit tests of the kernel building blocks through the C ABI (tcgen05 conv-GEMM, attention, the integer
length-regulator / duration kernels).  Compared with plain torch math on the same inputs or with the oracle."""
import ctypes as C
import math

import pytest
import torch

import _cases as CS
from daft_exprt_b200 import _lib, layouts
from oracle import daft_oracle as O

pytestmark = pytest.mark.gpu


def _lib_gpu():
    return _lib.load()


def _pack(w_gemm, nt, ks):
    lib = _lib_gpu()
    taps, K, N = w_gemm.shape
    n = lib.dx_packed_weight_elems(taps, K, N, nt, ks)
    out = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.dx_pack_weight_bf16(_lib.ptr(w_gemm), taps, K, N, nt, ks, _lib.ptr(out), _lib.stream_ptr()))
    return out


def _conv_gemm_ref(a, w, taps):
    """a: (M, K) bf16-rounded fp32, w: (taps, K, N) bf16-rounded fp32 → (M, N) in fp64."""
    M = a.shape[0]
    half = taps // 2
    a64, w64 = a.double(), w.double()
    out = torch.zeros(M, w.shape[2], dtype=torch.float64, device=a.device)
    for j in range(taps):
        sh = j - half
        src = torch.zeros_like(a64)
        if sh == 0:
            src = a64
        elif sh < 0:
            src[-sh:] = a64[:M + sh]
        else:
            src[:M - sh] = a64[sh:]
        out += src @ w64[j]
    return out


@pytest.mark.parametrize("M,K,N,taps,nt,ks", [
    (128, 64, 128, 1, 128, 64),       # single tile, single stage
    (128, 128, 128, 1, 128, 64),      # two K stages
    (300, 128, 384, 1, 128, 64),      # QKV shape, ragged M
    (300, 128, 128, 3, 128, 64),      # conv taps (shifted descriptors)
    (515, 128, 1024, 3, 256, 32),     # conv1 of the FFN
    (515, 1024, 128, 3, 128, 64),     # conv2 of the FFN
    (260, 96, 1024, 3, 256, 32),      # pre-net conv0 (padded mel)
    (260, 1024, 1024, 3, 256, 32),    # pre-net conv1
    (200, 128, 80, 1, 128, 64),       # mel projection (N padded to the tile)
])
def test_tc_gemm_matches_fp64(M, K, N, taps, nt, ks):
    lib = _lib_gpu()
    g = torch.Generator(device="cpu").manual_seed(M * 7 + K + N + taps)
    a = (torch.randn(M, K, generator=g)).to(torch.bfloat16).float().cuda()
    w = (torch.randn(taps, K, N, generator=g) / math.sqrt(K * taps)).to(torch.bfloat16).float().cuda()
    a_il = layouts.to_interleaved(a, cap=(M + 127) // 128 * 128)
    wp = _pack(w, nt, ks)
    out = torch.full((M, N), float("nan"), device="cuda")
    _lib.check(lib.dx_tc_gemm_test(_lib.ptr(a_il), a_il.shape[1], M, K, N, taps, _lib.ptr(wp), nt, ks, _lib.ptr(out),
                                   _lib.stream_ptr()), "dx_tc_gemm_test")
    torch.cuda.synchronize()
    ref = _conv_gemm_ref(a, w, taps)
    err = (out.double() - ref).abs().max().item()
    assert err <= 2e-3, f"max-abs {err}"      # fp32 accumulation of exact bf16 products


def _attention_ref(qkv_rows, starts, lengths):
    """qkv_rows: (total, 384) fp32 → (total, 128): per utterance, per head softmax(q k^T / 8) v."""
    out = torch.zeros(qkv_rows.shape[0], 128, dtype=torch.float64, device=qkv_rows.device)
    x = qkv_rows.double()
    for b, n in enumerate(lengths):
        rs = starts[b]
        for h in range(2):
            q = x[rs:rs + n, h * 64:(h + 1) * 64]
            k = x[rs:rs + n, 128 + h * 64:128 + (h + 1) * 64]
            v = x[rs:rs + n, 256 + h * 64:256 + (h + 1) * 64]
            s = torch.softmax(q @ k.t() * 0.125, dim=-1)
            out[rs:rs + n, h * 64:(h + 1) * 64] = s @ v
    return out


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
@pytest.mark.parametrize("lengths", [[40], [128], [129, 64, 7], [300, 255, 130, 1]])
def test_attention_matches_fp64(mode, lengths):
    lib = _lib_gpu()
    B, n_max = len(lengths), max(lengths)
    cap = lib.dx_row_cap(B, n_max)
    g = torch.Generator().manual_seed(sum(lengths))
    total = sum(lengths) + _lib.GAP * B
    qkv = torch.randn(total, 384, generator=g)
    if mode == "bf16":
        qkv = qkv.to(torch.bfloat16).float()
    qkv = qkv.cuda()
    starts, acc = [], 0
    for n in lengths:
        starts.append(acc)
        acc += n + _lib.GAP
    lens = torch.tensor(lengths, dtype=torch.long, device="cuda")
    scratch = torch.zeros((B + 1) + 2 * cap, dtype=torch.int32, device="cuda")
    if mode == "fp32":
        q_in = torch.zeros(cap, 384, device="cuda")
        q_in[:total] = qkv
        out = torch.zeros(cap, 128, device="cuda")
        _lib.check(lib.dx_test_attention(_lib.MODE_F32, _lib.ptr(q_in), _lib.ptr(out), 0, _lib.ptr(lens), B, n_max,
                                         _lib.ptr(scratch), cap, _lib.stream_ptr()), "dx_test_attention")
        torch.cuda.synchronize()
        got = out[:total].double()
        tol = 2e-5
    else:
        q_il = layouts.to_interleaved(qkv, cap=cap)
        o_il = torch.zeros(16, q_il.shape[1], 8, dtype=torch.bfloat16, device="cuda")
        _lib.check(lib.dx_test_attention(_lib.MODE_BF16, _lib.ptr(q_il), _lib.ptr(o_il), q_il.shape[1], _lib.ptr(lens), B,
                                         n_max, _lib.ptr(scratch), cap, _lib.stream_ptr()), "dx_test_attention")
        torch.cuda.synchronize()
        got = layouts.from_interleaved(o_il, total).double()
        tol = 3e-2        # P and the output are rounded to bf16
    ref = _attention_ref(qkv, starts, lengths)
    valid = torch.zeros(total, dtype=torch.bool, device="cuda")
    for b, n in enumerate(lengths):
        valid[starts[b]:starts[b] + n] = True
    err = (got - ref)[valid].abs().max().item()
    assert err <= tol, f"max-abs {err}"


def test_length_regulator_integers_bit_exact():
    lib = _lib_gpu()
    hp = CS.hparams()
    batch = CS.synthetic.forward_batch(hp, batch=9, lmin=3, lmax=77, seed=123)
    dint, in_len = batch[2], batch[5]
    ref = O.length_regulator_integers(dint, in_len)
    B, L = dint.shape
    T = ref["t_max"]
    d = dint.cuda()
    il = in_len.cuda()
    cs = torch.empty(B, L, dtype=torch.long, device="cuda")
    two_mu = torch.empty_like(cs)
    out_len = torch.empty(B, dtype=torch.long, device="cuda")
    f2s = torch.empty(B, T, dtype=torch.long, device="cuda")
    _lib.check(lib.dx_length_regulator_integers(_lib.ptr(d), _lib.ptr(il), B, L, T, _lib.ptr(cs), _lib.ptr(two_mu),
                                                _lib.ptr(out_len), _lib.ptr(f2s), _lib.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(cs.cpu(), ref["cumsum"])
    assert torch.equal(two_mu.cpu(), ref["two_mu"])
    assert torch.equal(out_len.cpu(), ref["output_lengths"])
    assert torch.equal(f2s.cpu(), ref["frame2symbol"])
    assert int(cs.max()) == T


def test_int_durations_bit_exact_and_error_rows():
    lib = _lib_gpu()
    hp = CS.hparams()
    g = torch.Generator().manual_seed(77)
    B, L = 48, 61
    hops = torch.rand(B, L, generator=g) * 7
    dur = hops * (hp.hop_length / hp.sampling_rate)
    dur[0, 3] = 0.001
    dur[1, -9:] = 0.0
    dur[2, :] = 0.0                 # the reference raises on an all-zero row
    dur[3, 5:] = 0.0
    d_dev = dur.clone().cuda().contiguous()
    dint = torch.empty(B, L, dtype=torch.long, device="cuda")
    status = torch.empty(B, dtype=torch.int32, device="cuda")
    _lib.check(lib.dx_int_durations(_lib.ptr(d_dev), B, L, hp.sampling_rate, hp.filter_length, hp.hop_length, 0,
                                    _lib.ptr(dint), _lib.ptr(status), _lib.stream_ptr()))
    torch.cuda.synchronize()
    status = status.cpu()
    n_ok = 0
    for b in range(B):
        row = dur[b:b + 1].clone()
        try:
            rd, ri = O.get_int_durations(row, hp)
            ok = True
        except Exception:
            ok = False
        assert (status[b].item() == 0) == ok, f"row {b}: status {status[b].item()} vs oracle ok={ok}"
        if ok:
            n_ok += 1
            assert torch.equal(dint[b].cpu(), ri[0]), f"row {b}"
            assert torch.equal(d_dev[b].cpu(), rd[0]), f"row {b}"
    assert n_ok >= B - 4


@pytest.mark.parametrize("pt", ["add", "multiply"])
def test_prosody_transform(pt):
    hp = CS.hparams()
    from daft_exprt_b200.model import DaftExprt
    model = DaftExprt(hp).cuda()
    g = torch.Generator().manual_seed(3)
    B, L = 5, 33
    pitch = torch.randn(B, L, generator=g)
    pitch[:, ::7] = 0.0
    fac = torch.randn(B, L, generator=g) * (20.0 if pt == "add" else 0.5)
    spk = torch.randint(0, 2, (B,), generator=g)
    if pt == "add":
        ref = O.pitch_shift(pitch.clone(), fac, hp, spk)
        got = model.pitch_shift(pitch.clone().cuda(), fac.cuda(), hp, spk.cuda())
    else:
        ref = O.pitch_multiply(pitch.clone(), fac)
        got = model.pitch_multiply(pitch.clone().cuda(), fac.cuda())
    torch.cuda.synchronize()
    got = got.cpu()
    finite = torch.isfinite(ref)
    assert torch.equal(torch.isfinite(got), finite)
    assert (got[finite] - ref[finite]).abs().max() <= 2e-5
