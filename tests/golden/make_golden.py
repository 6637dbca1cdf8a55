"""Generate tests/golden/*.npz from the LIVE reference (build container only).

    python tests/golden/make_golden.py

Imports ``/root/reference/src/daft_exprt/model.py`` through ``oracle/ref_harness.py`` (stubbed
librosa/matplotlib), loads the synthetic weights ``make_state_dict(hp, 1234)`` strictly into the
reference ``DaftExprt``, runs ``forward`` / ``inference`` / ``accent_encoder`` on the seeded batches
below and freezes the outputs.  Inputs are NOT stored: they are rebuilt from the same seeds by
``daft_exprt_b200.synthetic`` (torch CPU generators are deterministic for a given torch version; the
GPU box runs the same image).  A checksum of the weights and of every input tensor is stored so a
drifted generator is detected instead of silently comparing different problems.
"""
import os
import sys
import hashlib
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from daft_exprt_b200.hparams import default_hparams          # noqa: E402
from daft_exprt_b200.weights import make_state_dict           # noqa: E402
from daft_exprt_b200 import synthetic                         # noqa: E402
from oracle import ref_harness                                # noqa: E402

WEIGHT_SEED = 1234

FORWARD_CASES = {
    "fwd_tiny": dict(batch=3, lmin=5, lmax=12, seed=7),
    "fwd_pad": dict(batch=4, lmin=20, lmax=40, seed=3),
    "fwd_cfg1": dict(batch=1, lmin=60, lmax=60, seed=0, const_dur=5),
}
INFERENCE_CASES = {
    "inf_tiny_add": dict(batch=3, lmin=5, lmax=12, seed=11, pitch_transform="add"),
    "inf_tiny_mul": dict(batch=3, lmin=5, lmax=12, seed=11, pitch_transform="multiply"),
    "inf_pad_add": dict(batch=5, lmin=15, lmax=45, seed=13, pitch_transform="add"),
}


def tensor_digest(t):
    return hashlib.sha256(t.detach().contiguous().cpu().numpy().tobytes()).hexdigest()[:16]


def inputs_digest(tensors):
    return "|".join(tensor_digest(t) for t in tensors)


def weights_digest(sd):
    h = hashlib.sha256()
    for k, v in sd.items():
        h.update(k.encode())
        h.update(v.numpy().tobytes())
    return h.hexdigest()[:16]


def main():
    hp = default_hparams()
    sd = make_state_dict(hp, WEIGHT_SEED)
    ref = ref_harness.build_reference_model(hp, sd)
    out_dir = os.path.dirname(os.path.abspath(__file__))
    wd = weights_digest(sd)
    for name, kw in FORWARD_CASES.items():
        batch = synthetic.forward_batch(hp, **kw)
        with torch.no_grad():
            r = ref(tuple(t.clone() for t in batch))
            acc = ref.accent_encoder(batch[6].clone(), batch[7].clone(), batch[8].clone(), batch[9].clone())
        np.savez_compressed(
            os.path.join(out_dir, name + ".npz"),
            weights_digest=wd, inputs_digest=inputs_digest(batch),
            speaker_preds=r[0].numpy(), film_dec=r[1][3].numpy(), mel=r[3][0].numpy(),
            output_lengths=r[3][1].numpy(), weights=r[4].numpy(), accent_emb=acc.numpy())
        print(name, "mel", tuple(r[3][0].shape), "weights", tuple(r[4].shape))
    for name, kw in INFERENCE_CASES.items():
        kw = dict(kw)
        pt = kw["pitch_transform"]
        inputs, prosody, ext_emb, acc_emb = synthetic.inference_batch(hp, **kw)
        dig = inputs_digest(list(inputs) + [prosody[k] for k in sorted(prosody)] + [ext_emb, acc_emb])
        with torch.no_grad():
            enc, dec, w = ref.inference(tuple(t.clone() for t in inputs), pt, hp,
                                        {k: v.clone() for k, v in prosody.items()}, ext_emb.clone(),
                                        acc_emb.clone())
        np.savez_compressed(
            os.path.join(out_dir, name + ".npz"),
            weights_digest=wd, inputs_digest=dig,
            duration_preds=enc[0].numpy(), durations_int=enc[1].numpy(), energy_preds=enc[2].numpy(),
            pitch_preds=enc[3].numpy(), mel=dec[0].numpy(), output_lengths=dec[1].numpy(),
            weights=w.numpy())
        print(name, "mel", tuple(dec[0].shape))


if __name__ == "__main__":
    main()
