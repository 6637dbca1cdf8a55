"""TEST INFRASTRUCTURE ONLY — imports the *real* reference: from /root/reference (read-only, build container) or,
where that does not exist (the GPU box), from the byte-identical staged copy ``oracle/_ref`` that
``oracle/make_ref.py`` makes (git-ignored, travels with the gpurun snapshot).

It is used to (1) pin ``oracle/daft_oracle.py`` (our CPU restatement) against the live reference,
(2) generate the golden vectors under ``tests/golden/`` (``tests/golden/make_golden.py``), and
(3) time the reference's own CPU forward in ``bench.py --impl reference`` / ``cpu_baseline``.
Nothing in the product package may import it.

Accommodations (none modify the reference):
  * ``librosa`` / ``matplotlib`` are not installed here and ``daft_exprt/model.py:11`` pulls them
    in transitively → stub modules are injected into ``sys.modules`` before the import;
  * ``HyperParams`` is not instantiated (MFA file asserts, ``hparams.py:218-229``) → a plain
    namespace with the model-relevant fields is used instead.
"""
import os
import sys
import types

_LIVE_SRC = "/root/reference/src"
_STAGED_SRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _pick_source():
    # DX_REF_FORCE_STAGED=1: use the staged copy even where /root/reference exists (tests of the GPU-box situation)
    if os.environ.get("DX_REF_FORCE_STAGED") != "1" and os.path.isfile(os.path.join(_LIVE_SRC, "daft_exprt", "model.py")):
        return _LIVE_SRC, "/root/reference"
    if os.path.isfile(os.path.join(_STAGED_SRC, "daft_exprt", "model.py")):
        from oracle import make_ref
        if make_ref.verify():
            return _STAGED_SRC, "oracle/_ref (staged copy of the reference, sha256-verified)"
    return None, None


REFERENCE_SRC, REFERENCE_KIND = _pick_source()


def reference_available():
    return REFERENCE_SRC is not None


def _stub(name, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def import_reference_model():
    """Return the reference's ``daft_exprt.model`` module."""
    if not reference_available():
        raise RuntimeError("reference not present (neither /root/reference nor a verified oracle/_ref copy)")
    for name in ("librosa", "matplotlib", "soundfile"):
        try:
            __import__(name)
        except Exception:
            _stub(name, use=lambda *a, **k: None)
    try:
        import librosa.filters  # noqa: F401
    except Exception:
        lf = _stub("librosa.filters", mel=lambda *a, **k: None)
        sys.modules["librosa"].filters = lf
    try:
        import matplotlib.pyplot  # noqa: F401
    except Exception:
        pp = _stub("matplotlib.pyplot")
        sys.modules["matplotlib"].pyplot = pp
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import daft_exprt.model as ref_model
    return ref_model


def build_reference_model(hp, state_dict):
    """Instantiate the reference ``DaftExprt`` on CPU (eval) and load ``state_dict`` strictly."""
    import copy
    ref_model = import_reference_model()
    model = ref_model.DaftExprt(copy.deepcopy(hp))
    model.load_state_dict(state_dict, strict=True)
    model.eval()
    return model


def import_reference_generate():
    """Return the reference's ``daft_exprt.generate`` module (the synthesis driver).

    ``generate.py:17`` pulls in the text cleaners, whose third-party imports (``unidecode``, ``inflect``) are not installed
    here and are not on the tensor path → stubbed like librosa above."""
    import_reference_model()
    import warnings
    for name in ("unidecode", "inflect"):
        try:
            __import__(name)
        except Exception:
            m = _stub(name)
            m.__getattr__ = lambda attr: (lambda *a, **k: None)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")          # the reference's regex literals raise SyntaxWarning on 3.12
        import daft_exprt.generate as ref_generate
    return ref_generate


class _ZeroVocoder:
    """Stands in for HiFi-GAN (no weights offline): ``generate.py:327`` only needs ``infer``."""

    def infer(self, mel_spec):
        import numpy as np
        return np.zeros(mel_spec.shape[1] * 256, dtype=np.float32)


def run_reference_generate(model, hp, output_dir, batch=None, **kw):
    """Run the reference's ``generate_batch_mel_specs`` (``batch=True``) or ``generate_mel_specs`` on CPU.

    The reference moves every tensor with ``.cuda(gpu, non_blocking=True)`` (``generate.py:205-210,273-278``); there is no
    GPU here, so ``torch.Tensor.cuda`` is replaced by the identity for the duration of the call (the model is on the CPU).
    Plotting (``generate.py:324-326``, matplotlib absent) is replaced by a no-op and the vocoder by zeros; neither touches
    the returned predictions."""
    import torch
    G = import_reference_generate()
    saved_cuda, saved_plot = torch.Tensor.cuda, G.plot_2d_data
    torch.Tensor.cuda = lambda self, *a, **k: self
    G.plot_2d_data = lambda *a, **k: None
    try:
        with torch.no_grad():
            if batch:
                return G.generate_batch_mel_specs(model=model, output_dir=output_dir, hparams=hp, n_jobs=1,
                                                  vocoder=_ZeroVocoder(), **kw)
            return G.generate_mel_specs(model=model, output_dir=output_dir, hparams=hp, vocoder=_ZeroVocoder(), **kw)
    finally:
        torch.Tensor.cuda, G.plot_2d_data = saved_cuda, saved_plot


def run_reference_fine_tuning(hp, batches, predictions):
    """Run the reference's ``fine_tuning`` main loop (``fine_tune.py:22-129``) on the CPU with a stand-in model.

    ``batches``: what its data loader would yield (tuples whose items 11 / 12 are feature_dirs / feature_files);
    ``predictions[k] = (mel_spec_preds (B, n_mels, T_max) tensor, output_lengths (B,) tensor)`` for batch k.  Only the writer
    half is under test, so the model, the checkpoint and the loader are replaced; ``librosa.load`` (absent) is replaced by
    the product's scipy reader for files already at the target rate and ``librosa_mel_fn`` by a constant filter bank (the
    loop only uses the target mel's shape).  Files land in ``dirname(hp.training_files)/fine_tuning_dataset``."""
    import numpy as np
    import torch
    import_reference_model()
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import daft_exprt.fine_tune as FT
        import daft_exprt.extract_features as EF
    from daft_exprt_b200.fine_tune import default_wav_loader

    class _Model:
        def __init__(self, *_a, **_k):
            self.k = -1

        def cuda(self, *_a, **_k):
            return self

        def load_state_dict(self, *_a, **_k):
            return None

        def eval(self):
            return self

        def parse_batch(self, _gpu, batch):
            self.k += 1
            return batch, None

        def __call__(self, _inputs):
            return None, None, None, predictions[self.k], None

    saved = (FT.DaftExprt, FT.prepare_data_loaders, FT.librosa, torch.cuda.set_device, torch.load, EF.librosa_mel_fn)
    fake_librosa = types.SimpleNamespace(load=lambda path, sr=None: default_wav_loader(path, sr))
    FT.DaftExprt = _Model
    FT.prepare_data_loaders = lambda *_a, **_k: (list(batches), None, None, None)
    FT.librosa = fake_librosa
    torch.cuda.set_device = lambda *_a, **_k: None
    torch.load = lambda *_a, **_k: {"state_dict": {}}
    EF.librosa_mel_fn = lambda sr, n_fft, n_mels, fmin, fmax: np.ones((n_mels, 1 + n_fft // 2), dtype=np.float32)
    try:
        FT.fine_tuning(hp)
    finally:
        FT.DaftExprt, FT.prepare_data_loaders, FT.librosa, torch.cuda.set_device, torch.load, EF.librosa_mel_fn = saved
    return os.path.join(os.path.dirname(hp.training_files), "fine_tuning_dataset")
