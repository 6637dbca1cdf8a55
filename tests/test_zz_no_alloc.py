"""GPU: the allocation-free attention entry point (dx_attention_static: row map in a file-scope __device__ array of the
library, include/daft_b200.h) against the existing entry point (dx_test_attention, caller scratch), through the C ABI.
Collected last on purpose."""
import ctypes

import pytest
import torch

import _cases as CS  # noqa: F401  (sets sys.path)
from daft_exprt_b200 import _lib, layouts

pytestmark = pytest.mark.gpu


def _inputs(mode, lengths, seed):
    lib = _lib.load()
    B, n_max = len(lengths), max(lengths)
    cap = lib.dx_row_cap(B, n_max)
    total = sum(lengths) + _lib.GAP * B
    qkv = torch.randn(total, 384, generator=torch.Generator().manual_seed(seed))
    lens = torch.tensor(lengths, dtype=torch.long, device="cuda")
    if mode == _lib.MODE_F32:
        q = torch.zeros(cap, 384, device="cuda")
        q[:total] = qkv.cuda()
        return lib, B, n_max, cap, total, lens, q, 0, lambda: torch.zeros(cap, 128, device="cuda")
    q = layouts.to_interleaved(qkv.to(torch.bfloat16).float().cuda(), cap=cap)
    return lib, B, n_max, cap, total, lens, q, q.shape[1], lambda: torch.zeros(16, q.shape[1], 8, dtype=torch.bfloat16,
                                                                                 device="cuda")


def _valid_rows(out, mode, lengths, total):
    rows = out[:total] if mode == _lib.MODE_F32 else layouts.from_interleaved(out, total)
    keep, acc = [], 0
    for n in lengths:
        keep.append(rows[acc:acc + n])
        acc += n + _lib.GAP
    return torch.cat(keep).float()


@pytest.mark.parametrize("mode", [_lib.MODE_F32, _lib.MODE_BF16])
def test_static_scratch_attention_equals_caller_scratch_attention_twice(mode):
    # two calls in a row with DIFFERENT shapes: the second one finds the first one's row map in the shared array
    for seed, lengths in ((1, [300, 255, 130, 1]), (2, [129, 64, 7])):
        lib, B, n_max, cap, total, lens, q, rows_phys, new_out = _inputs(mode, lengths, seed)
        want, got = new_out(), new_out()
        scratch = torch.zeros((B + 1) + 2 * cap, dtype=torch.int32, device="cuda")
        _lib.check(lib.dx_test_attention(mode, _lib.ptr(q), _lib.ptr(want), rows_phys, _lib.ptr(lens), B, n_max,
                                         _lib.ptr(scratch), cap, _lib.stream_ptr()), "dx_test_attention")
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]
        _lib.check(lib.dx_attention_static(mode, _lib.ptr(q), _lib.ptr(got), rows_phys, _lib.ptr(lens), B, n_max,
                                           _lib.stream_ptr()), "dx_attention_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == before      # nothing went through torch either
        a, b = _valid_rows(want, mode, lengths, total), _valid_rows(got, mode, lengths, total)
        assert torch.isfinite(a).all() and a.abs().max() > 0
        # same kernels, same row map: bit-identical (the tolerance of the existing attention tests is 2e-5 / 3e-2)
        assert torch.equal(a, b), f"max-abs {(a - b).abs().max().item()}"


def test_static_scratch_attention_rejects_sizes_beyond_its_maximum():
    lib = _lib.load()
    assert lib.dx_attention_static_max_rows() == 32768
    lens = torch.tensor([10], dtype=torch.long, device="cuda")
    x = torch.zeros(8, device="cuda")
    too_long = 32768                       # 1 * (32768 + 3) rows > 32768
    st = lib.dx_attention_static(_lib.MODE_F32, _lib.ptr(x), _lib.ptr(x), 0, _lib.ptr(lens), 1, too_long, _lib.stream_ptr())
    assert st == 2 and b"exceeds the static row map" in lib.dx_last_error_string()
    st = lib.dx_attention_static(_lib.MODE_F32, _lib.ptr(x), _lib.ptr(x), 0, _lib.ptr(lens), 256, 8, _lib.stream_ptr())
    assert st == 2 and b"outside 1..255" in lib.dx_last_error_string()
    st = lib.dx_attention_static(_lib.MODE_F32, None, _lib.ptr(x), 0, _lib.ptr(lens), 1, 8, _lib.stream_ptr())
    assert st == 2 and isinstance(lib.dx_last_error_string(), (bytes, ctypes.c_char_p))
