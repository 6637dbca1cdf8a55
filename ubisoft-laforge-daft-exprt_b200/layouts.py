"""Host-side helpers for the device layouts of DESIGN.md (used by tests and tools, not by the hot path).

Channel-interleaved bf16 activations: ``[C/8][rows_phys][8]`` with ``rows_phys = HALO + cap + TAIL``; stream row
``r`` lives at physical row ``HALO + r``; halo / tail rows are zero.
"""
import torch

from . import _lib


def rows_phys(cap):
    return _lib.HALO + cap + _lib.TAIL


def to_interleaved(x, cap=None):
    """x: (rows, C) float tensor → bf16 (C/8, rows_phys, 8) on the same device."""
    rows, c = x.shape
    assert c % 8 == 0
    cap = rows if cap is None else cap
    out = torch.zeros(c // 8, rows_phys(cap), 8, dtype=torch.bfloat16, device=x.device)
    out[:, _lib.HALO:_lib.HALO + rows, :] = x.to(torch.bfloat16).view(rows, c // 8, 8).permute(1, 0, 2)
    return out.contiguous()


def from_interleaved(x_il, rows):
    """bf16 (C/8, rows_phys, 8) → float32 (rows, C)."""
    nch = x_il.shape[0]
    return x_il[:, _lib.HALO:_lib.HALO + rows, :].permute(1, 0, 2).reshape(rows, nch * 8).float()


def pack_rows(x_padded, lengths):
    """(B, N, C) padded tensor → packed row stream (total_rows, C) with DX_GAP zero rows after each utterance,
    plus the row_start list.  Mirrors dx_build_rows."""
    B, N, C = x_padded.shape
    pieces, starts, acc = [], [], 0
    for b in range(B):
        n = int(lengths[b])
        starts.append(acc)
        pieces.append(x_padded[b, :n])
        pieces.append(x_padded.new_zeros(_lib.GAP, C))
        acc += n + _lib.GAP
    starts.append(acc)
    return torch.cat(pieces, dim=0), starts
