"""Generate tests/golden/gen_*.npz from the LIVE reference's synthesis driver (build container only).

    python tests/golden/make_generate_golden.py

Runs the reference's ``generate_mel_specs`` (``/root/reference/src/daft_exprt/generate.py:333-399``) on CPU through
``oracle/ref_harness.run_reference_generate`` (``Tensor.cuda`` -> identity, plots off, zero vocoder) with the synthetic
weights of ``make_golden.py`` loaded into the reference ``DaftExprt``, on the seeded requests below, and freezes the
returned predictions.  The requests are rebuilt from their seeds by ``daft_exprt_b200.synthetic.synthesis_request``; a
digest of the request is stored so a drifted generator is detected.
"""
import copy
import hashlib
import json
import os
import sys
import tempfile
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from daft_exprt_b200.hparams import default_hparams          # noqa: E402
from daft_exprt_b200.weights import make_state_dict           # noqa: E402
from daft_exprt_b200 import synthetic                         # noqa: E402
from make_golden import WEIGHT_SEED                           # noqa: E402

SOURCE_STATS = {"energy": {"mean": 1.5, "std": 0.8}, "pitch": {"mean": 4.9, "std": 0.2}}
GENERATE_CASES = {
    # name: (synthesis_request kwargs, generate_mel_specs kwargs)
    "gen_add": (dict(n=7, lmin=4, lmax=14, seed=5, pitch_transform="add"),
                dict(batch_size=3, source_stats=None, alpha_dur=1.0, alpha_pitch=1.0, alpha_energy=1.0)),
    "gen_mul_src": (dict(n=6, lmin=6, lmax=20, seed=6, pitch_transform="multiply"),
                    dict(batch_size=4, source_stats=SOURCE_STATS, alpha_dur=1.3, alpha_pitch=0.7, alpha_energy=1.2)),
}
FIELDS = ("duration", "duration_int", "energy", "pitch", "mel", "weight")


def request_digest(req):
    h = hashlib.sha256()
    plain = {k: v for k, v in req.items() if not torch.is_tensor(v)}
    h.update(json.dumps(plain, sort_keys=True).encode())
    for k in sorted(k for k, v in req.items() if torch.is_tensor(v)):
        h.update(req[k].numpy().tobytes())
    return h.hexdigest()[:16]


def main():
    from oracle import ref_harness
    hp = default_hparams()
    ref = ref_harness.build_reference_model(hp, make_state_dict(hp, WEIGHT_SEED))
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, (rk, gk) in GENERATE_CASES.items():
        req = synthetic.synthesis_request(hp, **rk)
        with tempfile.TemporaryDirectory() as tmp:
            preds = ref_harness.run_reference_generate(ref, hp, tmp, **copy.deepcopy(req), **gk)
        arrays = {"request_digest": request_digest(req), "names": np.array(list(preds.keys()))}
        for k, vals in preds.items():
            for f, v in zip(FIELDS, vals):
                arrays[f"{k}/{f}"] = np.ascontiguousarray(v)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **arrays)
        print(name, len(preds), "utterances", sum(v[4].shape[1] for v in preds.values()), "frames")


if __name__ == "__main__":
    main()
