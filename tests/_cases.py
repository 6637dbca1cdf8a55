"""Shared helpers: rebuild the golden cases' inputs from their seeds and load the frozen outputs."""
import os
import sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

from daft_exprt_b200.hparams import default_hparams      # noqa: E402
from daft_exprt_b200.weights import make_state_dict       # noqa: E402
from daft_exprt_b200 import synthetic                     # noqa: E402

sys.path.insert(0, GOLDEN_DIR)
import make_golden as MG                                  # noqa: E402

_SD = {}


def hparams():
    return default_hparams()


def state_dict(seed=MG.WEIGHT_SEED):
    if seed not in _SD:
        _SD[seed] = make_state_dict(default_hparams(), seed)
    return _SD[seed]


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return {k: z[k] for k in z.files}


def forward_case(name):
    hp = default_hparams()
    batch = synthetic.forward_batch(hp, **MG.FORWARD_CASES[name])
    g = load_golden(name)
    assert str(g["inputs_digest"]) == MG.inputs_digest(batch), "synthetic generator drifted"
    assert str(g["weights_digest"]) == MG.weights_digest(state_dict()), "weight generator drifted"
    return hp, batch, g


def inference_case(name):
    hp = default_hparams()
    kw = dict(MG.INFERENCE_CASES[name])
    inputs, prosody, ext_emb, acc_emb = synthetic.inference_batch(hp, **kw)
    g = load_golden(name)
    dig = MG.inputs_digest(list(inputs) + [prosody[k] for k in sorted(prosody)] + [ext_emb, acc_emb])
    assert str(g["inputs_digest"]) == dig, "synthetic generator drifted"
    return hp, kw["pitch_transform"], inputs, prosody, ext_emb, acc_emb, g


def t(a):
    return torch.from_numpy(np.asarray(a))


_MEASURED = {}


def record(key, value):
    """Keep the measured error of a bf16 / fp32 comparison (worst value per key over the session) in
    gpurun_out/parity_measured.json, so that the tolerances in the tests can be read against what the hardware produced."""
    import json
    value = float(value)
    _MEASURED[key] = max(value, _MEASURED.get(key, value))
    try:
        d = os.path.join(ROOT, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, "parity_measured.json")
        old = json.load(open(path)) if os.path.isfile(path) else {}
        for k, v in _MEASURED.items():
            old[k] = max(v, old.get(k, v))
        json.dump(old, open(path, "w"), indent=1, sort_keys=True)
    except (OSError, ValueError):
        pass


# ---- architectures with other block counts / an accent_encoder sub-dict of its own (tests/golden/make_arch_golden.py) ----
import make_arch_golden as MA                             # noqa: E402


def arch_case(name):
    """-> (hparams, state dict, forward batch, inference tuple or None, golden dict)"""
    hp = MA.arch_hparams(name)
    key = ("arch", name)
    if key not in _SD:
        _SD[key] = make_state_dict(hp, MG.WEIGHT_SEED)
    sd = _SD[key]
    fkw, ikw = MA.ARCH_CASES[name][3:5]
    batch = synthetic.forward_batch(hp, **fkw)
    g = load_golden(name)
    assert str(g["inputs_digest"]) == MG.inputs_digest(batch), "synthetic generator drifted"
    assert str(g["weights_digest"]) == MG.weights_digest(sd), "weight generator drifted"
    inf = None
    if ikw is not None:
        kw = dict(ikw)
        inputs, prosody, ext_emb, acc_emb = synthetic.inference_batch(hp, **kw)
        dig = MG.inputs_digest(list(inputs) + [prosody[k] for k in sorted(prosody)] + [ext_emb, acc_emb])
        assert str(g["inf_inputs_digest"]) == dig, "synthetic generator drifted"
        inf = (kw["pitch_transform"], inputs, prosody, ext_emb, acc_emb)
    return hp, sd, batch, inf, g
