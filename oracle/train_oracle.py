"""TEST INFRASTRUCTURE ONLY — the reference's training iteration (train.py:421-434) on the CPU as the gradient oracle.

Runs the UNMODIFIED reference classes (``oracle/ref_harness.py``: /root/reference in the build container, the staged
``oracle/_ref`` copy on the GPU box): ``DaftExprt.forward`` (model.py:889-948) under autograd - in eval mode (dropout =
identity, the parity configuration of SURVEY.md section 8d cfg5), or in train mode with the dropout MASKS injected (the random
draw of ``torch.nn.functional.dropout`` is replaced by masks the caller supplies, in call order; everything else, including
which tensors get dropped and the 1 / (1 - p) rescale, is the reference's) -, ``DaftExprtLoss`` (loss.py:57-159, pitch-consistency off) and
``loss.backward()``.  Only tests/ and bench.py's cpu_baseline leg may import this file.
"""
import copy

import torch

from oracle import ref_harness as R


class _InjectedDropout:
    """Stands in for ``torch.nn.functional.dropout`` while the reference's train-mode forward runs: call k multiplies its
    input by ``masks[k] / (1 - p)``.  Shapes and probabilities are checked, so a call-order mismatch fails loudly."""

    def __init__(self, masks):
        self.masks = list(masks)
        self.k = 0

    def __call__(self, input, p=0.5, training=True, inplace=False):
        if not training or p == 0.0:
            return input
        assert self.k < len(self.masks), f"dropout call {self.k}: the reference makes more dropout calls than masks were given"
        m, pm = self.masks[self.k]
        assert tuple(m.shape) == tuple(input.shape), f"dropout call {self.k}: input {tuple(input.shape)} vs mask {tuple(m.shape)}"
        assert abs(pm - p) < 1e-7, f"dropout call {self.k}: p={p} vs mask p={pm}"
        self.k += 1
        return input * (m.to(input.dtype) * (1.0 / (1.0 - p)))


def reference_step(hp, state_dict, batch, iteration, num_threads=None, dropout_masks=None):
    """-> (loss float, individual-loss dict, {parameter name: gradient tensor}) for the 12-tuple ``batch`` (CPU tensors).
    ``dropout_masks``: None = eval mode; else the list of (bool keep-mask, p) for ``model.train()``, in call order."""
    if num_threads:
        torch.set_num_threads(num_threads)
    model = R.build_reference_model(hp, state_dict)          # eval()
    if dropout_masks is not None:
        import torch.nn.functional as F
        model.train()
        injected = _InjectedDropout(dropout_masks)
        stock = F.dropout
        F.dropout = injected
        try:
            return _step(model, hp, batch, iteration, injected)
        finally:
            F.dropout = stock
    return _step(model, hp, batch, iteration, None)


def _step(model, hp, batch, iteration, injected):
    import daft_exprt.loss as ref_loss
    crit = ref_loss.DaftExprtLoss("cpu", copy.deepcopy(hp))
    inputs = tuple(t.clone() for t in batch)
    targets = (inputs[1], inputs[3], inputs[4], inputs[8], inputs[9], inputs[10])      # model.py:884-885
    for p in model.parameters():
        p.grad = None
    outputs = model(inputs)
    loss, individual = crit(outputs, targets, iteration)
    if injected is not None:
        assert injected.k == len(injected.masks), f"the reference made {injected.k} dropout calls, {len(injected.masks)} masks given"
    loss.backward()
    grads = {n: (p.grad.detach().clone() if p.grad is not None else torch.zeros_like(p)) for n, p in model.named_parameters()}
    return float(loss.detach()), individual, grads
