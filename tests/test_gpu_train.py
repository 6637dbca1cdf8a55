"""GPU: the training step of BASELINE config 5 (dx_train_step / dx_adam_step through train.TrainEngine) against the
reference's own iteration run on the CPU (oracle/train_oracle.py: the unmodified reference classes under autograd).

Gates (fp32 mode): loss terms within 1e-5 relative; every one of the 184 gradients within 1e-3 max-abs AND 2e-3 relative
L1 of the reference's; an Adam step equal to torch.optim.Adam's on the same gradients; bit-identical repeat."""
import pytest
import torch

import _cases as C
from daft_exprt_b200.train import TrainEngine, loss_dict, update_learning_rate
from oracle import ref_harness as R
from oracle import train_oracle as TO

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not R.reference_available(), reason="reference (oracle/_ref) not present")]

GRAD_MAX_ABS = 1e-3
GRAD_REL_L1 = 2e-3
LOSS_REL = 1e-5


def _check_step(batch, iteration, hp=None):
    hp = hp or C.hparams()
    sd = C.state_dict()
    ref_loss, ref_ind, ref_grads = TO.reference_step(hp, sd, batch, iteration)
    eng = TrainEngine(hp, sd)
    losses = loss_dict(eng.forward_backward(tuple(t.cuda() for t in batch), iteration=iteration))
    assert abs(losses["loss"] - ref_loss) <= LOSS_REL * abs(ref_loss), (losses["loss"], ref_loss)
    for k, v in ref_ind.items():
        assert abs(losses[k] - v) <= LOSS_REL * max(abs(v), 1e-3), (k, losses[k], v)
    grads = eng.grads_state_dict()
    assert sorted(grads) == sorted(ref_grads) and len(grads) == 184
    worst_abs = worst_rel = 0.0
    for name, r in ref_grads.items():
        o = grads[name].cpu()
        assert o.shape == r.shape, name
        d = (o - r).abs()
        rel = d.sum().item() / max(r.abs().sum().item(), 1e-12)
        worst_abs, worst_rel = max(worst_abs, d.max().item()), max(worst_rel, rel)
        assert d.max().item() <= GRAD_MAX_ABS, f"{name}: max-abs {d.max().item():.3e}"
        assert rel <= GRAD_REL_L1 or r.abs().sum().item() < 1e-6, f"{name}: rel-L1 {rel:.3e}"
    return eng, worst_abs, worst_rel


@pytest.mark.parametrize("kw,iteration", [
    (dict(batch=4, lmin=5, lmax=30, seed=17), 1234),          # ragged, warm-up phase of the adversarial weight
    (dict(batch=3, lmin=12, lmax=12, seed=18, const_dur=3), 20000),   # no padding anywhere (true zero padding at the ends)
    (dict(batch=2, lmin=1, lmax=2, seed=19), 1),               # 1- and 2-symbol utterances
])
def test_gradients_match_reference_autograd(kw, iteration):
    batch = C.synthetic.forward_batch(C.hparams(), **kw)
    _, worst_abs, worst_rel = _check_step(batch, iteration)
    C.record(f"train_grad_max_abs_B{kw['batch']}", worst_abs)
    C.record(f"train_grad_rel_l1_B{kw['batch']}", worst_rel)


def test_gradients_match_reference_autograd_cfg5_size():
    """BASELINE config 5's per-GPU batch: 32 utterances drawn as cfg2 (20-150 symbols, 1-9 frames per symbol)."""
    batch = C.synthetic.forward_batch(C.hparams(), batch=32, lmin=20, lmax=150, seed=42)
    _, worst_abs, worst_rel = _check_step(batch, 5000)
    C.record("train_grad_max_abs_B32", worst_abs)
    C.record("train_grad_rel_l1_B32", worst_rel)


def test_no_post_multipliers_and_no_energy_term():
    """post_mult_weight = 0 removes the (2, 8) parameter (model.py:760-761, loss.py:92-95); energy_consistency_weight = 0
    removes that term (loss.py:113)."""
    hp = C.hparams()
    hp.post_mult_weight = 0.0
    hp.energy_consistency_weight = 0.0
    sd = {k: v for k, v in C.state_dict().items() if k != "style_adapter.post_multipliers"}
    batch = C.synthetic.forward_batch(hp, batch=3, lmin=4, lmax=20, seed=23)
    ref_loss, ref_ind, ref_grads = TO.reference_step(hp, sd, batch, 77)
    eng = TrainEngine(hp, sd)
    losses = loss_dict(eng.forward_backward(tuple(t.cuda() for t in batch), iteration=77))
    assert abs(losses["loss"] - ref_loss) <= LOSS_REL * abs(ref_loss)
    assert losses["energy_consistency_loss"] == 0.0 and losses["post_mult_loss"] == 0.0
    grads = eng.grads_state_dict()
    assert len(grads) == 183
    for name, r in ref_grads.items():
        assert (grads[name].cpu() - r).abs().max().item() <= GRAD_MAX_ABS, name


def test_step_is_deterministic_and_adam_equals_torch():
    hp = C.hparams()
    sd = C.state_dict()
    batch = tuple(t.cuda() for t in C.synthetic.forward_batch(hp, batch=4, lmin=5, lmax=30, seed=17))
    eng = TrainEngine(hp, sd)
    eng.forward_backward(batch, iteration=10)
    g1 = eng.grads.clone()
    eng.forward_backward(batch, iteration=10)
    assert torch.equal(g1, eng.grads), "two identical steps must give bit-identical gradients (no atomics)"
    # three Adam steps on the flat buffers against torch.optim.Adam with the reference's settings (train.py:278-280)
    p_ref = eng.params.clone().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], betas=tuple(hp.betas), eps=hp.epsilon, weight_decay=hp.weight_decay, amsgrad=False)
    for it in (1, 2, 3):
        lr = update_learning_rate(hp, it)
        eng.forward_backward(batch, iteration=it)
        for grp in opt.param_groups:
            grp["lr"] = lr
        p_ref.grad = eng.grads.clone()
        opt.step()
        eng.adam(lr)
        assert (eng.params - p_ref.detach()).abs().max().item() <= 2.5e-7      # 1-2 ulp at |w| ~ 1
    # the updated weights still export in checkpoint layout and load into the inference mirror
    from daft_exprt_b200.model import DaftExprt
    m = DaftExprt(hp, mode="fp32").cuda()
    m.load_state_dict(eng.state_dict())
    losses_fwd = loss_dict(eng.forward_backward(batch, iteration=4, forward_only=True)).copy()
    _, mel = eng.forward_backward(batch, iteration=4, forward_only=True, want_mel=True)
    assert torch.equal(m(batch)[3][0], mel), "training forward and DX_MODE_F32 inference forward are the same kernels"
    assert losses_fwd["loss"] > 0


def test_loss_goes_down():
    """Ten iterations on one batch at a fixed learning rate: the loss the step reports must fall (end-to-end sanity of
    gradient signs, layouts and the optimiser)."""
    hp = C.hparams()
    batch = tuple(t.cuda() for t in C.synthetic.forward_batch(hp, batch=8, lmin=10, lmax=40, seed=29))
    eng = TrainEngine(hp, C.state_dict())
    first = None
    for it in range(1, 11):
        v = float(eng.step(batch, iteration=it, lr=2e-4)[0])
        first = v if first is None else first
    assert v < 0.97 * first, (first, v)        # measured: 4.51 -> 4.18


def test_step_does_not_depend_on_what_the_workspace_held():
    """The workspace is caller memory with arbitrary contents: a step on a workspace full of NaN bit patterns gives the
    same loss and bit-identical gradients as on a zeroed one (0 x stale-NaN would poison a weight gradient; found on
    out_proj.weight, whose operand rows between utterances were never written)."""
    hp = C.hparams()
    batch = tuple(t.cuda() for t in C.synthetic.forward_batch(hp, batch=8, lmin=10, lmax=40, seed=29))
    got = []
    for fill in (0x00, 0xFF):
        eng = TrainEngine(hp, C.state_dict())
        eng._workspace(batch[0].shape[0], batch[0].shape[1], batch[8].shape[2]).fill_(fill)
        losses = eng.forward_backward(batch, iteration=1).clone()
        torch.cuda.synchronize()
        got.append((losses, eng.grads.clone()))
    assert torch.isfinite(got[1][1]).all()
    assert torch.equal(got[0][0], got[1][0]) and torch.equal(got[0][1], got[1][1])


# ---- tensor-core GEMM mode (TrainEngine(gemm="bf16"), DX_TRAIN_TC_GEMM) --------------------------------------------------
# bf16 operands with fp32 accumulation for the conv / linear forward, dgrad and wgrad contractions; everything else fp32.
# Gates set from the measured errors (profiles/r02_parity_measured.json: train_tc_*): the loss within 5e-3 relative of
# the reference's; every one of the 184 gradients within 2.5e-1 relative L1 of the reference's autograd and their median
# within 3e-2 (bf16 rounding of activations and gradients accumulates through the ~30 contractions of the backward chain:
# the deepest tensors - the accent pre-net's, and the speaker classifier's behind the gradient reversal, whose gradients are
# tiny - carry the largest error: 9e-2 at B = 4, 1.6e-1 at B = 32).
TC_LOSS_REL = 5e-3
TC_GRAD_REL_L1_MAX = 2.5e-1
TC_GRAD_REL_L1_MEDIAN = 3e-2


@pytest.mark.parametrize("kw,iteration,tag", [
    (dict(batch=4, lmin=5, lmax=30, seed=17), 1234, "B4"),
    (dict(batch=32, lmin=20, lmax=150, seed=42), 5000, "B32"),
])
def test_tensor_core_mode_gradients_against_reference_autograd(kw, iteration, tag):
    hp = C.hparams()
    sd = C.state_dict()
    batch = C.synthetic.forward_batch(hp, **kw)
    ref_loss, _, ref_grads = TO.reference_step(hp, sd, batch, iteration)
    eng = TrainEngine(hp, sd, gemm="bf16")
    losses = loss_dict(eng.forward_backward(tuple(t.cuda() for t in batch), iteration=iteration))
    assert abs(losses["loss"] - ref_loss) <= TC_LOSS_REL * abs(ref_loss), (losses["loss"], ref_loss)
    grads = eng.grads_state_dict()
    rels = []
    for name, r in ref_grads.items():
        o = grads[name].cpu()
        assert torch.isfinite(o).all(), name
        if r.abs().sum().item() < 1e-6:
            continue
        rel = (o - r).abs().sum().item() / r.abs().sum().item()
        rels.append(rel)
        assert rel <= TC_GRAD_REL_L1_MAX, f"{name}: rel-L1 {rel:.3e}"
    rels.sort()
    C.record(f"train_tc_grad_rel_l1_max_{tag}", rels[-1])
    C.record(f"train_tc_grad_rel_l1_median_{tag}", rels[len(rels) // 2])
    C.record(f"train_tc_loss_rel_{tag}", abs(losses["loss"] - ref_loss) / abs(ref_loss))
    assert rels[len(rels) // 2] <= TC_GRAD_REL_L1_MEDIAN, rels[len(rels) // 2]


def test_tensor_core_mode_is_deterministic_workspace_independent_and_learns():
    hp = C.hparams()
    batch = tuple(t.cuda() for t in C.synthetic.forward_batch(hp, batch=8, lmin=10, lmax=40, seed=29))
    got = []
    for fill in (0x00, 0xFF, 0xFF):
        eng = TrainEngine(hp, C.state_dict(), gemm="bf16")
        eng._workspace(batch[0].shape[0], batch[0].shape[1], batch[8].shape[2]).fill_(fill)
        losses = eng.forward_backward(batch, iteration=1).clone()
        torch.cuda.synchronize()
        got.append((losses, eng.grads.clone()))
    assert torch.isfinite(got[1][1]).all()
    for other in got[1:]:
        assert torch.equal(got[0][0], other[0]) and torch.equal(got[0][1], other[1])
    eng = TrainEngine(hp, C.state_dict(), gemm="bf16")
    first = None
    for it in range(1, 11):
        v = float(eng.step(batch, iteration=it, lr=2e-4)[0])
        first = v if first is None else first
    assert v < 0.97 * first, (first, v)


def test_accumulation_and_gradient_clipping_follow_the_reference_iteration():
    """train.py:421-445: loss / accumulation_steps per micro-batch, summed gradients, clip_grad_norm_, Adam.  Checked
    against torch's own clip_grad_norm_ + Adam applied to the engine's per-micro-batch gradients."""
    hp = C.hparams()
    hp.accumulation_steps = 2
    hp.grad_clip_thresh = 0.05
    sd = C.state_dict()
    mbs = [tuple(t.cuda() for t in C.synthetic.forward_batch(hp, batch=3, lmin=5, lmax=25, seed=s)) for s in (71, 72)]
    eng = TrainEngine(hp, sd)
    g = []
    for mb in mbs:
        eng.forward_backward(mb, iteration=7, loss_scale=0.5)
        g.append(eng.grads.clone())
    p_ref = eng.params.clone().requires_grad_(True)
    p_ref.grad = g[0] + g[1]
    ref_norm = torch.nn.utils.clip_grad_norm_([p_ref], hp.grad_clip_thresh)
    assert float(ref_norm) > hp.grad_clip_thresh                  # the clip is active in this case
    opt = torch.optim.Adam([p_ref], lr=1e-4, betas=tuple(hp.betas), eps=hp.epsilon, weight_decay=hp.weight_decay)
    opt.step()
    eng2 = TrainEngine(hp, sd)
    total, norm = eng2.iterate(mbs, iteration=7, lr=1e-4)
    assert abs(float(norm) - float(ref_norm)) <= 1e-5 * float(ref_norm)
    assert (eng2.params - p_ref.detach()).abs().max().item() <= 2.5e-7
    with pytest.raises(ValueError):
        eng2.iterate(mbs[:1], iteration=8)


# ---- train-mode forward: dropout (model.py:165-168,189,216,656-670) -----------------------------------------------------
# The engine draws its masks from a counter-based generator; dx_dropout_mask exports them in the reference's padded
# coordinates and the oracle injects them into the UNMODIFIED reference model in train() mode (oracle/train_oracle.py), so
# both sides drop exactly the same elements.  Same gates as the eval-mode tensor-core step.

def _dropout_case(hp, kw, iteration, seed, p=None):
    if p is not None:
        for name in ("phoneme_encoder", "frame_decoder", "accent_encoder"):
            if hasattr(hp, name):
                d = dict(getattr(hp, name))
                d["attn_dropout"], d["conv_dropout"] = p
                setattr(hp, name, d)
    sd = C.state_dict()
    batch = C.synthetic.forward_batch(hp, **kw)
    eng = TrainEngine(hp, sd, gemm="bf16", train_mode=True, seed=seed)
    B, L = batch[0].shape
    T = batch[8].shape[2]
    masks = [(m.cpu(), pm) for m, pm in eng.dropout_masks(B, L, T, iteration=iteration)]
    losses = loss_dict(eng.forward_backward(tuple(t.cuda() for t in batch), iteration=iteration))
    return eng, batch, sd, masks, losses


def _rel_l1(grads, ref_grads):
    rels = {}
    for name, r in ref_grads.items():
        o = grads[name].cpu()
        assert torch.isfinite(o).all(), name
        if r.abs().sum().item() < 1e-6:
            continue
        rels[name] = (o - r).abs().sum().item() / r.abs().sum().item()
    return rels


# worst-tensor gate 0.4 (eval mode: TC_GRAD_REL_L1_MAX = 0.25): the worst tensors are one- to 128-element biases whose
# gradient is a sum with heavy cancellation (measured: the range predictor's scalar bias 0.26 at B = 8, a speaker-classifier
# bias behind the gradient reversal 0.29 at B = 3 with p = 0.3); dropout's 1 / (1 - p) rescale widens the bf16 rounding of
# every branch it touches.  The median gate is the eval-mode one, the loss gate too (profiles/r02_parity_measured.json).
TC_DROPOUT_GRAD_REL_L1_MAX = 0.4


@pytest.mark.parametrize("kw,iteration,p,tag", [
    (dict(batch=4, lmin=5, lmax=30, seed=17), 1234, None, "B4"),              # hparams' own 0.1 / 0.1
    (dict(batch=3, lmin=8, lmax=24, seed=31), 300, (0.3, 0.2), "B3_p03"),     # heavier dropout: a wrong mask cannot hide
    (dict(batch=8, lmin=20, lmax=150, seed=43), 5000, None, "B8"),            # several attention tiles per utterance
])
def test_train_mode_dropout_against_reference_with_the_same_masks(kw, iteration, p, tag):
    hp = C.hparams()
    eng, batch, sd, masks, losses = _dropout_case(hp, kw, iteration, seed=2024, p=p)
    assert len(masks) == 39
    for m, pm in masks:                         # keep rate of every site is 1 - p within 5 sigma
        n = m.numel()
        assert abs(m.float().mean().item() - (1 - pm)) <= 5 * (pm * (1 - pm) / n) ** 0.5 + 1e-6
    ref_loss, _, ref_grads = TO.reference_step(hp, sd, batch, iteration, dropout_masks=masks)
    eval_loss, _, _ = TO.reference_step(hp, sd, batch, iteration)
    assert abs(ref_loss - eval_loss) > 2 * TC_LOSS_REL * abs(ref_loss), "the case must tell train mode from eval mode"
    assert abs(losses["loss"] - ref_loss) <= TC_LOSS_REL * abs(ref_loss), (losses["loss"], ref_loss, eval_loss)
    rels = sorted(_rel_l1(eng.grads_state_dict(), ref_grads).items(), key=lambda kv: kv[1])
    C.record(f"train_dropout_grad_rel_l1_max_{tag}", rels[-1][1])
    C.record(f"train_dropout_grad_rel_l1_median_{tag}", rels[len(rels) // 2][1])
    C.record(f"train_dropout_loss_rel_{tag}", abs(losses["loss"] - ref_loss) / abs(ref_loss))
    assert rels[-1][1] <= TC_DROPOUT_GRAD_REL_L1_MAX, rels[-1]
    assert rels[len(rels) // 2][1] <= TC_GRAD_REL_L1_MEDIAN, rels[len(rels) // 2]


def test_train_mode_masks_of_another_seed_do_not_match():
    """Control for the test above: against the reference run with the masks of a DIFFERENT seed the gradients are far
    outside the gate, so the gate really checks that both sides dropped the same elements."""
    hp = C.hparams()
    kw, iteration = dict(batch=3, lmin=8, lmax=24, seed=31), 300
    eng, batch, sd, masks, _ = _dropout_case(hp, kw, iteration, seed=1)
    other = TrainEngine(hp, sd, gemm="bf16", train_mode=True, seed=2)
    B, L = batch[0].shape
    wrong = [(m.cpu(), pm) for m, pm in other.dropout_masks(B, L, batch[8].shape[2], iteration=iteration)]
    assert any(not torch.equal(a[0], b[0]) for a, b in zip(masks, wrong))
    _, _, g_right = TO.reference_step(hp, sd, batch, iteration, dropout_masks=masks)
    _, _, g_wrong = TO.reference_step(hp, sd, batch, iteration, dropout_masks=wrong)
    right = sorted(_rel_l1(eng.grads_state_dict(), g_right).values())
    wrong_r = sorted(_rel_l1(eng.grads_state_dict(), g_wrong).values())
    assert right[len(right) // 2] <= TC_GRAD_REL_L1_MEDIAN
    assert wrong_r[len(wrong_r) // 2] > 5 * right[len(right) // 2], (right[len(right) // 2], wrong_r[len(wrong_r) // 2])


def test_train_mode_seeding_determinism_and_zero_probability():
    hp = C.hparams()
    sd = C.state_dict()
    batch = tuple(t.cuda() for t in C.synthetic.forward_batch(hp, batch=4, lmin=5, lmax=30, seed=17))
    eng = TrainEngine(hp, sd, gemm="bf16", train_mode=True, seed=7)
    eng.forward_backward(batch, iteration=3)
    g3 = eng.grads.clone()
    eng.forward_backward(batch, iteration=4)
    g4 = eng.grads.clone()
    eng.forward_backward(batch, iteration=3)
    assert torch.equal(g3, eng.grads), "same (seed, iteration): same masks, bit-identical gradients"
    assert not torch.equal(g3, g4), "another iteration draws other masks"
    eng.forward_backward(batch, iteration=3, micro_batch=1)
    assert not torch.equal(g3, eng.grads), "micro-batches of one iteration draw different masks"
    # validation (forward_only) is eval mode whatever the engine's mode (train.py:155-209)
    ev = TrainEngine(hp, sd, gemm="bf16")
    a = eng.forward_backward(batch, iteration=3, forward_only=True).clone()
    b = ev.forward_backward(batch, iteration=3, forward_only=True).clone()
    assert torch.equal(a, b)
    # p = 0 everywhere in train mode is the eval-mode step, bit for bit
    hp0 = C.hparams()
    for name in ("phoneme_encoder", "frame_decoder", "accent_encoder"):
        if hasattr(hp0, name):
            d = dict(getattr(hp0, name))
            d["attn_dropout"] = d["conv_dropout"] = 0.0
            setattr(hp0, name, d)
    z = TrainEngine(hp0, sd, gemm="bf16", train_mode=True, seed=7)
    z.forward_backward(batch, iteration=3)
    ev.forward_backward(batch, iteration=3)
    assert torch.equal(z.grads, ev.grads)
    # dropout needs the tensor-core path; the fp32 path refuses instead of silently skipping it
    with pytest.raises(ValueError):
        TrainEngine(hp, sd, gemm="fp32", train_mode=True)


def test_train_mode_learns():
    hp = C.hparams()
    batch = tuple(t.cuda() for t in C.synthetic.forward_batch(hp, batch=8, lmin=10, lmax=40, seed=29))
    eng = TrainEngine(hp, C.state_dict(), gemm="bf16", train_mode=True, seed=11)
    for it in range(1, 21):
        eng.step(batch, iteration=it, lr=2e-4)
    first = float(TrainEngine(hp, C.state_dict(), gemm="bf16").forward_backward(batch, iteration=1, forward_only=True)[0])
    ev = TrainEngine(hp, eng.state_dict(), gemm="bf16")
    last = float(ev.forward_backward(batch, iteration=1, forward_only=True)[0])
    assert last < 0.97 * first, (first, last)
