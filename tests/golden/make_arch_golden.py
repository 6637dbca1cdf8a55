"""Generate tests/golden/arch_*.npz from the LIVE reference (build container only): architectures whose
``hparams.accent_encoder`` sub-dict differs from ``phoneme_encoder`` (model.py:627-631) and whose stacks have other block
counts than the default 4 / 4 / 4.

    python tests/golden/make_arch_golden.py

Same recipe as make_golden.py: synthetic weights ``make_state_dict(hp, 1234)`` loaded strictly into the reference
``DaftExprt`` built from the modified hparams, seeded synthetic batches, outputs frozen with digests of weights and inputs.
"""
import os
import sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from daft_exprt_b200.hparams import default_hparams          # noqa: E402
from daft_exprt_b200.weights import make_state_dict           # noqa: E402
from daft_exprt_b200 import synthetic                         # noqa: E402
import make_golden as MG                                      # noqa: E402

# name -> (accent encoder blocks, phoneme encoder blocks, frame decoder blocks, forward batch, inference batch or None)
ARCH_CASES = {
    "arch_232": (2, 3, 2, dict(batch=4, lmin=10, lmax=40, seed=21), dict(batch=3, lmin=8, lmax=30, seed=22, pitch_transform="add")),
    "arch_143": (1, 4, 3, dict(batch=3, lmin=5, lmax=25, seed=23), None),
    "arch_411": (4, 1, 1, dict(batch=2, lmin=30, lmax=60, seed=24), None),
}


def arch_hparams(name):
    n_ae, n_pe, n_fd = ARCH_CASES[name][:3]
    hp = default_hparams()
    hp.phoneme_encoder = dict(hp.phoneme_encoder, nb_blocks=n_pe)
    hp.frame_decoder = dict(hp.frame_decoder, nb_blocks=n_fd)
    # its own sub-dict, as real checkpoints carry it: other depth and other dropout rates than the phoneme encoder's
    hp.accent_encoder = dict(hp.phoneme_encoder, nb_blocks=n_ae, attn_dropout=0.2, conv_dropout=0.05)
    return hp


def main():
    from oracle import ref_harness
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, (n_ae, n_pe, n_fd, fkw, ikw) in ARCH_CASES.items():
        hp = arch_hparams(name)
        sd = make_state_dict(hp, MG.WEIGHT_SEED)
        ref = ref_harness.build_reference_model(hp, sd)
        assert len(ref.accent_encoder.blocks) == n_ae and len(ref.phoneme_encoder.blocks) == n_pe
        batch = synthetic.forward_batch(hp, **fkw)
        with torch.no_grad():
            r = ref(tuple(t.clone() for t in batch))
            acc = ref.accent_encoder(batch[6].clone(), batch[7].clone(), batch[8].clone(), batch[9].clone())
        out = dict(weights_digest=MG.weights_digest(sd), inputs_digest=MG.inputs_digest(batch),
                   speaker_preds=r[0].numpy(), film_dec=r[1][3].numpy(), mel=r[3][0].numpy(),
                   output_lengths=r[3][1].numpy(), weights=r[4].numpy(), accent_emb=acc.numpy())
        if ikw is not None:
            kw = dict(ikw)
            pt = kw["pitch_transform"]
            inputs, prosody, ext_emb, acc_emb = synthetic.inference_batch(hp, **kw)
            out["inf_inputs_digest"] = MG.inputs_digest(list(inputs) + [prosody[k] for k in sorted(prosody)] + [ext_emb, acc_emb])
            with torch.no_grad():
                enc, dec, w = ref.inference(tuple(t.clone() for t in inputs), pt, hp, {k: v.clone() for k, v in prosody.items()},
                                            ext_emb.clone(), acc_emb.clone())
            out.update(inf_durations_int=enc[1].numpy(), inf_mel=dec[0].numpy(), inf_output_lengths=dec[1].numpy(),
                       inf_weights=w.numpy())
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **out)
        print(name, "film_dec", tuple(r[1][3].shape), "mel", tuple(r[3][0].shape))


if __name__ == "__main__":
    main()
