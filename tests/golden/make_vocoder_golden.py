"""Generate tests/golden/voc_*.npz from the LIVE reference HiFi-GAN generator (build container only).

    python tests/golden/make_vocoder_golden.py

Loads ``/root/reference/src/daft_exprt/vocoder/hifigan.py`` by path (it needs only torch / numpy), builds the
reference ``HiFiGANGenerator(DEFAULT_CONFIG)``, loads the seeded weight-normed checkpoint
``oracle.hifigan_oracle.make_state_dict(seed=4321)`` strictly, calls ``remove_weight_norm()`` and vocodes every
utterance of each case ALONE on its cropped mel — the reference's ``HiFiGanVocoder.infer`` contract
(hifigan.py:211-226).  Inputs are rebuilt from their seeds by ``vocoder_case``; digests of weights and inputs are
stored so that a drifted generator is detected.
"""
import hashlib
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import hifigan_oracle as H          # noqa: E402

REF_FILE = "/root/reference/src/daft_exprt/vocoder/hifigan.py"
WEIGHT_SEED = 4321
VOCODER_CASES = {
    "voc_one": dict(lengths=[24], seed=21),
    "voc_ragged": dict(lengths=[17, 9, 1, 12], seed=22),
}


def vocoder_case(lengths, seed):
    """(mels (B, 80, T_max) zero-padded, lengths (B,) int64) from a seeded CPU generator."""
    g = torch.Generator().manual_seed(seed)
    B, T = len(lengths), max(lengths)
    mels = torch.zeros(B, 80, T)
    for b, n in enumerate(lengths):
        mels[b, :, :n] = torch.randn(80, n, generator=g) * 1.5 - 1.0
    return mels, torch.tensor(lengths, dtype=torch.int64)


def digest(tensors):
    h = hashlib.sha256()
    for t in tensors:
        h.update(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes())
    return h.hexdigest()[:16]


def load_reference_module():
    spec = importlib.util.spec_from_file_location("ref_hifigan", REF_FILE)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def reference_generator(sd):
    m = load_reference_module()
    gen = m.HiFiGANGenerator(m.DEFAULT_CONFIG)
    gen.load_state_dict(sd)
    gen.remove_weight_norm()
    return gen.eval()


def main():
    sd = H.make_state_dict(seed=WEIGHT_SEED)
    gen = reference_generator(sd)
    for name, kw in VOCODER_CASES.items():
        mels, lengths = vocoder_case(**kw)
        B, _, T = mels.shape
        audio = np.zeros((B, T * 256), np.float32)
        with torch.no_grad():
            for b in range(B):
                n = int(lengths[b])
                audio[b, : n * 256] = gen(mels[b : b + 1, :, :n]).squeeze().numpy()
        np.savez_compressed(os.path.join(os.path.dirname(__file__), name + ".npz"), audio=audio,
                            weights_digest=digest([sd[k] for k in sorted(sd)]), inputs_digest=digest([mels, lengths]))
        print(name, audio.shape, "std", audio.std(), "max", np.abs(audio).max())


if __name__ == "__main__":
    main()
