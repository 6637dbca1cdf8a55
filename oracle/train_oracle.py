"""TEST INFRASTRUCTURE ONLY — the reference's training iteration (train.py:421-434) on the CPU as the gradient oracle.

Runs the UNMODIFIED reference classes (``oracle/ref_harness.py``: /root/reference in the build container, the staged
``oracle/_ref`` copy on the GPU box): ``DaftExprt.forward`` (model.py:889-948) in eval mode (dropout = identity, the parity
configuration of SURVEY.md section 8d cfg5) under autograd, ``DaftExprtLoss`` (loss.py:57-159, pitch-consistency off) and
``loss.backward()``.  Only tests/ and bench.py's cpu_baseline leg may import this file.
"""
import copy

import torch

from oracle import ref_harness as R


def reference_step(hp, state_dict, batch, iteration, num_threads=None):
    """-> (loss float, individual-loss dict, {parameter name: gradient tensor}) for the 12-tuple ``batch`` (CPU tensors)."""
    if num_threads:
        torch.set_num_threads(num_threads)
    model = R.build_reference_model(hp, state_dict)          # eval()
    import daft_exprt.loss as ref_loss
    crit = ref_loss.DaftExprtLoss("cpu", copy.deepcopy(hp))
    inputs = tuple(t.clone() for t in batch)
    targets = (inputs[1], inputs[3], inputs[4], inputs[8], inputs[9], inputs[10])      # model.py:884-885
    for p in model.parameters():
        p.grad = None
    outputs = model(inputs)
    loss, individual = crit(outputs, targets, iteration)
    loss.backward()
    grads = {n: (p.grad.detach().clone() if p.grad is not None else torch.zeros_like(p)) for n, p in model.named_parameters()}
    return float(loss.detach()), individual, grads
