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


def test_ctypes_structs_match_the_header_compiled_as_c(tmp_path):
    """The header is plain C: compile it with gcc, print sizeof / offsetof of every struct field the bindings declare, and
    compare with the ctypes structures field by field (a mismatch would make every pointer after it garbage on the device)."""
    import ctypes
    import shutil
    import subprocess
    import pytest
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("gcc not available")
    structs = {"DxBlockWeights": _lib.DxBlockWeights, "DxModelWeights": _lib.DxModelWeights, "DxForwardArgs": _lib.DxForwardArgs,
               "DxTrainArgs": _lib.DxTrainArgs, "DxVcConv": _lib.DxVcConv, "DxVocoderWeights": _lib.DxVocoderWeights,
               "DxVocoderArgs": _lib.DxVocoderArgs}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "daft_b200.h"', 'int main(void) {']
    for name, cls in structs.items():
        lines.append(f'  printf("{name} %zu\\n", sizeof({name}));')
        for field, _t in cls._fields_:
            lines.append(f'  printf("{name}.{field} %zu\\n", offsetof({name}, {field}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Werror", "-I", os.path.dirname(HEADER), str(src), "-o", str(exe)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr       # also proves every ctypes field name exists in the header
    got = dict(line.split() for line in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for name, cls in structs.items():
        assert int(got[name]) == ctypes.sizeof(cls), (name, got[name], ctypes.sizeof(cls))
        for field, _t in cls._fields_:
            assert int(got[f"{name}.{field}"]) == getattr(cls, field).offset, (name, field)
    # and no field of the header is missing from the bindings: struct sizes equal, fields contiguous up to padding


def test_ctypes_signatures_have_the_headers_argument_counts_and_kinds():
    """Every prototype of the header against its ctypes signature: same number of arguments, pointers bound as pointers,
    64-bit integers as 64-bit, floats as floats (a wrong width silently shifts every later register / stack slot)."""
    import ctypes
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    protos = re.findall(r"\b([A-Za-z_][A-Za-z0-9_ \*]*?)\b(dx_[a-z0-9_]+)\s*\(([^()]*)\)\s*;", src)
    assert len(protos) == len(_lib.SIGNATURES)

    def kind(decl):
        decl = decl.strip()
        if decl in ("void", ""):
            return None
        if "*" in decl:
            return "ptr"
        base = decl.rsplit(None, 1)[0] if " " in decl else decl
        if "int64_t" in base or "uint64_t" in base:
            return "i64"
        if "float" in base:
            return "f32"
        return "i32"

    def ckind(t):
        if t in (ctypes.c_void_p, ctypes.c_char_p) or isinstance(t, type(ctypes.POINTER(ctypes.c_int))) and issubclass(t, ctypes._Pointer):
            return "ptr"
        if t in (ctypes.c_int64, ctypes.c_uint64):
            return "i64"
        if t is ctypes.c_float:
            return "f32"
        return "i32"

    for ret, name, args in protos:
        want = [k for k in (kind(a) for a in args.split(",")) if k is not None]
        restype, argtypes = _lib.SIGNATURES[name]
        assert [ckind(t) for t in argtypes] == want, (name, want, [ckind(t) for t in argtypes])
        rk = "ptr" if "*" in ret else ("i64" if "int64_t" in ret else "i32")
        assert ckind(restype) == rk, (name, ret, restype)


def test_weight_cache_key_sees_updates_and_replaced_parameters():
    """model._param_versions() (the key of the packed-weight cache, read on every call) without walking the module tree:
    it must change on in-place updates, on load_state_dict and when a Parameter OBJECT is replaced."""
    import torch
    from daft_exprt_b200.model import DaftExprt
    m = DaftExprt(C.hparams())
    k0 = m._param_versions()
    assert len(k0) == len(list(m.parameters())) == 184 and m._param_versions() == k0
    with torch.no_grad():
        m.frame_decoder.projection.linear_layer.bias.add_(1.0)                       # what an optimizer step does
    k1 = m._param_versions()
    assert k1 != k0
    m.load_state_dict(C.state_dict())
    k2 = m._param_versions()
    assert k2 != k1
    holder = m.frame_decoder.projection.linear_layer
    holder.bias = torch.nn.Parameter(holder.bias.detach().clone())                   # a new object with a fresh version counter
    k3 = m._param_versions()
    assert k3 != k2 and len(k3) == 184
    assert [id(p) for p in m.parameters()] == [i for i, _ in k3]                      # same order and objects as parameters()
