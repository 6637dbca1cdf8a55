"""CPU: the C-ABI library builds, loads, and exports every symbol include/daft_b200.h declares.
No compute calls (no GPU here)."""
import os
import re

import _cases as C  # noqa: F401  (sets sys.path)
from daft_exprt_b200 import _lib

HEADER = os.path.join(C.ROOT, "include", "daft_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dx_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    lib = _lib.load()
    names = declared_functions()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"libdaft_b200.so does not export {n}"
        assert n in _lib.SIGNATURES, f"ctypes binding missing for {n}"
    assert sorted(_lib.SIGNATURES) == names


def test_host_only_queries():
    lib = _lib.load()
    assert lib.dx_version() >= 100
    assert lib.dx_row_cap(64, 784) % 256 == 0 and lib.dx_row_cap(64, 784) >= 64 * 787
    assert lib.dx_forward_workspace_bytes(64, 150, 784, _lib.MODE_BF16) > 0
    assert lib.dx_forward_workspace_bytes(64, 150, 784, _lib.MODE_F32) > 0
    assert lib.dx_forward_workspace_bytes(0, 1, 1, 0) == -1
    assert b"plan" in lib.dx_last_error_string()
    assert lib.dx_packed_weight_elems(3, 128, 1024, 256, 32) == 3 * 128 * 1024
    assert lib.dx_packed_weight_elems(1, 128, 80, 128, 64) == 128 * 128     # N padded to the tile


def test_struct_sizes_match_header():
    # 18 pointers per block; the model struct is pointers + 3 block arrays + 8 int32 (sizes, block counts)
    import ctypes
    assert ctypes.sizeof(_lib.DxBlockWeights) == 18 * 8
    n_ptr = 19 + 14 + 8 + 4
    assert ctypes.sizeof(_lib.DxModelWeights) == n_ptr * 8 + 3 * 4 * 18 * 8 + 32
    assert ctypes.sizeof(_lib.DxForwardArgs) == 16 + 21 * 8 + 8 + 8


def test_model_mirror_has_reference_layout_and_fails_loudly_on_cpu():
    import pytest
    import torch
    from daft_exprt_b200.model import DaftExprt
    hp = C.hparams()
    m = DaftExprt(hp)
    sd = C.state_dict()
    assert list(m.state_dict().keys()) == list(sd.keys())
    m.load_state_dict({"module." + k: v for k, v in sd.items()})     # DDP prefix is stripped
    for k, v in m.state_dict().items():
        assert torch.equal(v, sd[k])
    batch = C.synthetic.forward_batch(hp, batch=2, lmin=3, lmax=5, seed=1)
    with pytest.raises(ValueError):
        m(batch[:11])
    if not torch.cuda.is_available():
        with pytest.raises(_lib.DxError):
            m(batch)           # no CPU fallback
