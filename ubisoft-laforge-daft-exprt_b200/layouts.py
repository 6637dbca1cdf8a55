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


# ----------------------------------------------------------------------------------------------------------------
# packed frame-side I/O (include/daft_b200.h DX_IO_PACKED_*; DaftExprt.forward(..., packed_io=(T_max, total)))
# ----------------------------------------------------------------------------------------------------------------
def frame_starts(output_lengths):
    """Exclusive prefix sum of the frame counts: utterance b's frames start at ``fs[b]``; ``fs[B]`` = total."""
    fs = [0]
    for n in output_lengths.tolist() if hasattr(output_lengths, "tolist") else output_lengths:
        fs.append(fs[-1] + int(n))
    return fs


def pack_frames(frames_energy, frames_pitch, mel_specs, output_lengths):
    """Padded (B, T) / (B, T) / (B, 80, T) → packed float[ΣT], float[ΣT], float[80·ΣT]: utterance b's mel is the
    ``(80, T_b)`` row-major block at ``80 · fs[b]`` (the layout of the reference's per-utterance ``.npy`` features,
    ``data_loader.py:150-160``).  Host-side helper for tests and callers that start from padded tensors; a loader that
    has the per-utterance arrays copies them straight into the packed staging block instead."""
    lens = [int(n) for n in output_lengths.tolist()]
    e = torch.cat([frames_energy[b, :n] for b, n in enumerate(lens)])
    p = torch.cat([frames_pitch[b, :n] for b, n in enumerate(lens)])
    m = torch.cat([mel_specs[b, :, :n].reshape(-1) for b, n in enumerate(lens)])
    return e.contiguous(), p.contiguous(), m.contiguous()


def mel_views(packed_mel, output_lengths, n_mel=80):
    """Per-utterance ``(80, T_b)`` views of a packed mel (no copies) — the crop ``fine_tune.py:79`` takes."""
    fs = frame_starts(output_lengths)
    return [packed_mel[n_mel * fs[b]:n_mel * fs[b + 1]].view(n_mel, fs[b + 1] - fs[b]) for b in range(len(fs) - 1)]


def unpack_mel(packed_mel, output_lengths, T_max=None, n_mel=80):
    """Packed mel → the reference's padded (B, 80, T_max) tensor with zeros behind every utterance."""
    views = mel_views(packed_mel, output_lengths, n_mel)
    T_max = max(v.shape[1] for v in views) if T_max is None else T_max
    out = packed_mel.new_zeros(len(views), n_mel, T_max)
    for b, v in enumerate(views):
        out[b, :, :v.shape[1]] = v
    return out
