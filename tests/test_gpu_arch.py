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
    hp = C.hparams()
    hp.accent_encoder = dict(hp.phoneme_encoder, attn_nb_heads=4)
    with pytest.raises(NotImplementedError):
        TrainEngine(hp, C.state_dict())


@pytest.mark.parametrize("name", ["arch_232", "arch_143"])
def test_training_step_on_other_block_counts(name):
    """Gradients of every parameter against the reference's autograd (fp32 mode: the 1e-3 gates of test_gpu_train.py;
    tensor-core mode: its gates), and the train-mode step with the engine's dropout masks injected into the reference - the
    accent encoder of these cases has its own dropout rates (0.2 / 0.05)."""
    import test_gpu_train as T
    from daft_exprt_b200.train import TrainEngine, loss_dict
    from oracle import ref_harness as R
    from oracle import train_oracle as TO
    if not R.reference_available():
        pytest.skip("reference (oracle/_ref) not present")
    hp, sd, batch, _, _ = C.arch_case(name)
    dev = tuple(t.cuda() for t in batch)
    ref_loss, ref_ind, ref_grads = TO.reference_step(hp, sd, batch, 3000)
    eng = TrainEngine(hp, sd)
    losses = loss_dict(eng.forward_backward(dev, iteration=3000))
    assert abs(losses["loss"] - ref_loss) <= T.LOSS_REL * abs(ref_loss)
    for k, v in ref_ind.items():
        assert abs(losses[k] - v) <= T.LOSS_REL * max(abs(v), 1e-3), (k, losses[k], v)
    grads = eng.grads_state_dict()
    assert sorted(grads) == sorted(ref_grads)
    # Gates of test_gpu_train.py, except that ONE ReLU kink event is tolerated: in arch_143 the reference's pre-activation of
    # the second pre-net conv is -4.5e-7 at (utterance 1, frame 0, channel 260) - within rounding of zero - and the engine's
    # summation order lands on the other side, so that one position's gradient passes the ReLU on one side only.  It shows as
    # outliers in that output channel of the conv's weight / bias gradient (the 4/4/4 model on the same batch has the same
    # event, tools/arch_grad_debug.py) and as a 2-3e-3 relative-L1 knock-on in the two tensors upstream of it.
    with_outliers = []
    for n, r in ref_grads.items():
        d = (grads[n].cpu() - r).abs()
        rel = d.sum().item() / max(r.abs().sum().item(), 1e-12)
        assert rel <= 5e-3 or r.abs().sum().item() < 1e-6, f"{n}: rel-L1 {rel:.3e}"
        n_out = int((d > T.GRAD_MAX_ABS).sum())
        if n_out:
            with_outliers.append(n)
            assert n_out <= max(1, int(5e-3 * d.numel())), f"{n}: {n_out} of {d.numel()} elements off by more than {T.GRAD_MAX_ABS}"
    assert len(with_outliers) <= 4 and all(k.startswith("accent_encoder.convs.") for k in with_outliers), with_outliers
    # tensor-core mode, train mode: same masks on both sides
    eng = TrainEngine(hp, sd, gemm="bf16", train_mode=True, seed=5)
    B, L = batch[0].shape
    masks = [(m.cpu(), pm) for m, pm in eng.dropout_masks(B, L, batch[8].shape[2], iteration=3000)]
    n_ae, n_pe, n_fd = C.MA.ARCH_CASES[name][:3]
    assert len(masks) == 3 + 3 * (n_ae + n_pe + n_fd)
    assert masks[3][1] == pytest.approx(0.2) and masks[5][1] == pytest.approx(0.05)      # the accent encoder's own rates
    losses = loss_dict(eng.forward_backward(dev, iteration=3000))
    ref_loss, _, ref_grads = TO.reference_step(hp, sd, batch, 3000, dropout_masks=masks)
    rels = sorted(T._rel_l1(eng.grads_state_dict(), ref_grads).values())
    C.record(f"train_dropout_grad_rel_l1_median_{name}", rels[len(rels) // 2])
    C.record(f"train_dropout_loss_rel_{name}", abs(losses["loss"] - ref_loss) / abs(ref_loss))
    if name == "arch_232":
        assert abs(losses["loss"] - ref_loss) <= T.TC_LOSS_REL * abs(ref_loss), (losses["loss"], ref_loss)
        assert rels[-1] <= T.TC_DROPOUT_GRAD_REL_L1_MAX and rels[len(rels) // 2] <= T.TC_GRAD_REL_L1_MEDIAN, (rels[-1], rels[len(rels) // 2])
    else:
        # arch_143 (one accent block, 3 utterances, loss 8.1 at these random weights) is ill-conditioned for bf16 operands
        # with or without dropout: the EVAL-mode tensor-core step is already 1.0e-2 off in the loss and 8e-2 in the median
        # gradient (tools/arch_grad_debug.py; the 4/4/4 model on the same batch: 1.4e-3 / 1.4e-2), so the masks cannot be
        # judged by the tight gates here - they are by arch_232 above and the cases of test_gpu_train.py.  Wide gates:
        assert abs(losses["loss"] - ref_loss) <= 2e-2 * abs(ref_loss), (losses["loss"], ref_loss)
        assert rels[len(rels) // 2] <= 0.15, rels[len(rels) // 2]
