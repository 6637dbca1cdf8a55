"""TEST / BASELINE INFRASTRUCTURE ONLY — recipe that stages the UNMODIFIED reference package under ``oracle/_ref/``.

    python oracle/make_ref.py            # build container only: needs /root/reference

The reference (claussss/ubisoft-laforge-daft-exprt) is pure Python; it cannot be pip-installed offline (librosa,
MFA asserts) and ``/root/reference`` does not exist on the GPU box.  This script copies the
reference's own ``src/daft_exprt`` package, byte for byte, from where it lies into ``oracle/_ref/daft_exprt`` so
that the live reference class travels to the GPU box with the gpurun snapshot (``oracle/_ref/`` is git-ignored — no
reference source enters this repository's history — but not gpurun-ignored).  ``bench.py --impl reference`` then times
the reference's own ``DaftExprt.forward`` (``cpu_baseline.kind = "reference"``), and the ``-m gpu`` tests can pin the
oracle port against it on the box as well.  Nothing under ``ubisoft-laforge-daft-exprt_b200/`` may import it.

``__graft_entry__.build()`` calls ``stage()`` whenever ``/root/reference`` is present.  A ``MANIFEST.json`` with the
sha256 of every staged file is written next to the copy so that a stale or edited copy is detectable.
"""
import hashlib
import json
import os
import shutil

REFERENCE_PKG = "/root/reference/src/daft_exprt"
HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(HERE, "_ref")
REF_PKG = os.path.join(REF_ROOT, "daft_exprt")


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 20), b""):
            h.update(chunk)
    return h.hexdigest()


def _wanted(rel):
    # the Python package + the REAPER binary its __init__ chmods (src/daft_exprt/__init__.py:17-18)
    return rel.endswith(".py") or rel == os.path.join("bin", "reaper", "linux", "reaper")


def stage(force=False):
    """Copy the reference package into oracle/_ref (no edits).  Returns the manifest dict, or None when the reference is
    not present (GPU box: the prebuilt copy is used as is)."""
    if not os.path.isfile(os.path.join(REFERENCE_PKG, "model.py")):
        return None
    manifest = {}
    for dirpath, _, files in os.walk(REFERENCE_PKG):
        for fn in files:
            src = os.path.join(dirpath, fn)
            rel = os.path.relpath(src, REFERENCE_PKG)
            if not _wanted(rel):
                continue
            dst = os.path.join(REF_PKG, rel)
            digest = _sha(src)
            manifest[rel] = digest
            if not force and os.path.isfile(dst) and _sha(dst) == digest:
                continue
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            shutil.copyfile(src, dst)
            shutil.copymode(src, dst)
    with open(os.path.join(REF_ROOT, "MANIFEST.json"), "w") as f:
        json.dump({"source": REFERENCE_PKG, "files": manifest}, f, indent=1, sort_keys=True)
    return manifest


def verify():
    """True when oracle/_ref holds every file of its manifest with the recorded digest."""
    mf = os.path.join(REF_ROOT, "MANIFEST.json")
    if not os.path.isfile(mf):
        return False
    files = json.load(open(mf))["files"]
    return all(os.path.isfile(os.path.join(REF_PKG, rel)) and _sha(os.path.join(REF_PKG, rel)) == d for rel, d in files.items())


if __name__ == "__main__":
    m = stage(force=True)
    print("reference not present" if m is None else f"staged {len(m)} files into {REF_PKG}; verify: {verify()}")
