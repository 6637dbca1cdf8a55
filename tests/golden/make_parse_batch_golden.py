"""Generate tests/golden/parse_batch.npz from the LIVE reference's ``DaftExprt.parse_batch`` (model.py:858-887).

    python tests/golden/make_parse_batch_golden.py      # build container only

The 14-tuple a loader hands over is rebuilt from seeds by ``parse_batch_case()`` below with deliberately non-canonical
dtypes (int32 symbols, float64 durations, float16 energies, int16 lengths ...) so that the dtype casts of
``parse_batch`` are what is being pinned; the reference's 12 inputs + 6 targets (values and dtypes) are frozen.
"""
import os
import sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from daft_exprt_b200.hparams import default_hparams          # noqa: E402
from daft_exprt_b200.weights import make_state_dict           # noqa: E402
from daft_exprt_b200 import synthetic                         # noqa: E402

NAMES_IN = ("symbols", "durations_float", "durations_int", "symbols_energy", "symbols_pitch", "input_lengths",
            "frames_energy", "frames_pitch", "mel_specs", "output_lengths", "speaker_ids", "spk_embs")
NAMES_TGT = ("durations_float", "symbols_energy", "symbols_pitch", "mel_specs", "output_lengths", "speaker_ids")


def parse_batch_case():
    """The 14-tuple (loader order, data_loader.py:261-263) with off-contract dtypes."""
    hp = default_hparams()
    b = synthetic.forward_batch(hp, batch=5, lmin=4, lmax=17, seed=77)
    dirs = [f"/feat/spk{int(s)}" for s in b[10]]
    files = [f"utt{i:02d}" for i in range(5)]
    batch14 = (b[0].to(torch.int32), b[1].to(torch.float64), b[2].to(torch.int32), b[3].to(torch.float16), b[4].to(torch.float64),
               b[5].to(torch.int16), b[6].to(torch.float64), b[7].to(torch.float16), b[8].to(torch.float64), b[9].to(torch.int32),
               b[10].to(torch.uint8), dirs, files, b[11].to(torch.float64))
    return hp, batch14


def main():
    from oracle import ref_harness
    hp, batch14 = parse_batch_case()
    ref = ref_harness.build_reference_model(hp, make_state_dict(hp, 1234))
    inputs, targets = ref.parse_batch("cpu", batch14)
    out = {}
    for n, t in zip(NAMES_IN, inputs):
        out["in_" + n] = t.numpy()
    for n, t in zip(NAMES_TGT, targets):
        out["tgt_" + n] = t.numpy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "parse_batch.npz")
    np.savez_compressed(path, **out)
    print(path, {k: (v.dtype, v.shape) for k, v in out.items()})


if __name__ == "__main__":
    main()
