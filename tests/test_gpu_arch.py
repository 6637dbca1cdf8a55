"""GPU: architectures other than the reference's default - 1 to 4 FFT blocks per stack, each stack its own count, and an
``hparams.accent_encoder`` sub-dict that differs from ``phoneme_encoder`` (model.py:627-631) - in both modes against golden
vectors frozen from the live reference built from the same hparams (tests/golden/make_arch_golden.py).  Same gates as
test_gpu_parity.py."""
import pytest
import torch

import _cases as C
import test_gpu_parity as P
from daft_exprt_b200.model import DaftExprt

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["fp32", "bf16"])
@pytest.mark.parametrize("name", sorted(C.MA.ARCH_CASES))
def test_forward_and_inference_vs_golden_on_other_block_counts(mode, name):
    hp, sd, batch, inf, g = C.arch_case(name)
    n_fd = C.MA.ARCH_CASES[name][2]
    m = DaftExprt(hp, mode=mode).cuda()
    m.load_state_dict(sd)
    dev = tuple(t.cuda() for t in batch)
    spk, film, enc, dec, w = m(dev)
    torch.cuda.synchronize()
    assert torch.equal(dec[1].cpu(), C.t(g["output_lengths"]))
    P._check_mel(dec[0], C.t(g["mel"]), batch[9], mode)
    P._check_weights(w, C.t(g["weights"]), batch[5], batch[9], mode)
    tol = 1e-3 if mode == "fp32" else P.FILM_BF16_MAX_ABS
    assert film[3].shape == (batch[0].shape[0], n_fd, 256)
    assert (film[3].cpu() - C.t(g["film_dec"])).abs().max().item() <= tol
    assert (spk.cpu() - C.t(g["speaker_preds"])).abs().max().item() <= tol
    acc = m.accent_encoder(dev[6], dev[7], dev[8], dev[9]).cpu()
    assert (acc - C.t(g["accent_emb"])).abs().max().item() <= (1e-4 if mode == "fp32" else P.ACCENT_BF16_MAX_ABS)
    if inf is not None:
        pt, inputs, prosody, ext_emb, acc_emb = inf
        enc, dec, w = m.inference(tuple(t.cuda() for t in inputs), pt, hp, {k: v.clone().cuda() for k, v in prosody.items()},
                                  ext_emb.cuda(), acc_emb.cuda())
        torch.cuda.synchronize()
        assert torch.equal(enc[1].cpu(), C.t(g["inf_durations_int"]))
        assert torch.equal(dec[1].cpu(), C.t(g["inf_output_lengths"]))
        P._check_mel(dec[0], C.t(g["inf_mel"]), dec[1], mode)


def test_block_kernel_variants_agree_on_an_odd_block_count():
    """3 phoneme-encoder blocks: the fused two-launch-per-block chain ends in the other ping-pong buffer than with 4; the
    four-launch path (debug knob 9) must give the same bits."""
    from daft_exprt_b200 import _lib
    hp, sd, batch, _, _ = C.arch_case("arch_232")
    m = DaftExprt(hp, mode="bf16").cuda()
    m.load_state_dict(sd)
    dev = tuple(t.cuda() for t in batch)
    a = m(dev)[3][0].clone()
    lib = _lib.load()
    lib.dx_set_debug(9, 1)
    try:
        b = m(dev)[3][0].clone()
    finally:
        lib.dx_set_debug(9, 0)
    assert torch.equal(a, b)


def test_what_is_still_refused():
    from daft_exprt_b200.train import TrainEngine
    hp = C.hparams()
    hp.phoneme_encoder = dict(hp.phoneme_encoder, nb_blocks=5)
    with pytest.raises(NotImplementedError):
        DaftExprt(hp)
    hp = C.hparams()
    hp.frame_decoder = dict(hp.frame_decoder, conv_channels=512)
    with pytest.raises(NotImplementedError):
        DaftExprt(hp)
    hp, sd, _, _, _ = C.arch_case("arch_143")
    with pytest.raises(NotImplementedError):
        TrainEngine(hp, sd)
