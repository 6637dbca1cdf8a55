"""CPU: host logic of the training step (BASELINE config 5): engine layout <-> checkpoint layout, the learning-rate and
adversarial-weight schedules against the reference's own functions, and the data-parallel gradient exchange over ``gloo``
with world_size 2 (the path's only collective)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _cases as C
from daft_exprt_b200 import train as T
from daft_exprt_b200.weights import state_dict_spec
from oracle import ref_harness as R


def test_field_map_covers_the_checkpoint_and_layouts_round_trip():
    hp = C.hparams()
    spec = state_dict_spec(hp)
    fmap = T.field_map()
    assert sorted(fmap) == sorted(n for n, _ in spec) and len(fmap) == 184
    g = torch.Generator().manual_seed(0)
    for name, shape in spec:
        t = torch.randn(*shape, generator=g)
        tr = fmap[name][1]
        e = T.to_engine(t, tr)
        assert e.is_contiguous() and e.numel() == t.numel()
        assert torch.equal(T.from_engine(e.reshape(-1), tuple(shape), tr), t), name
    # conv weights: engine [tap][Cin][Cout]; projections: [in][out]
    w = torch.arange(2 * 3 * 5.0).view(2, 3, 5)               # [Cout=2][Cin=3][k=5]
    assert T.to_engine(w, "conv").shape == (5, 3, 2) and T.to_engine(w, "conv")[4, 1, 0] == w[0, 1, 4]
    assert T.to_engine(w[0], "lin_t")[4, 2] == w[0, 2, 4]


@pytest.mark.skipif(not R.reference_available(), reason="reference not present")
def test_schedules_equal_the_reference():
    R.import_reference_model()
    import daft_exprt.loss as ref_loss
    hp = C.hparams()
    crit = ref_loss.DaftExprtLoss("cpu", hp)
    for it in (1, 17, 9999, 10000, 10001, 250000):
        assert T.adversarial_weight(hp, it) == crit.update_adversarial_weight(it)
    # update_learning_rate lives in train.py, whose imports (tensorboard, PIL) are absent here: restated formula of
    # train.py:148-160, checked at the warm-up boundary and for continuity
    assert T.update_learning_rate(hp, 0) == hp.initial_learning_rate
    assert abs(T.update_learning_rate(hp, hp.warmup_steps) - hp.max_learning_rate) < 1e-12
    assert T.update_learning_rate(hp, 4 * hp.warmup_steps) == pytest.approx(hp.max_learning_rate / 2)
    assert T.update_learning_rate(hp, 5000) == pytest.approx((hp.max_learning_rate - hp.initial_learning_rate) / 2 + hp.initial_learning_rate)


class _Flat:
    """Stand-in with the engine's flat gradient buffer (the exchange only touches ``self.grads``)."""

    def __init__(self, grads):
        self.grads = grads


def _ddp_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(100 + rank)
        local = torch.randn(1000, generator=g)
        eng = _Flat(local.clone())
        n = T.TrainEngine.all_reduce_grads(eng)
        # every rank now holds the SUM; the Adam kernel applies 1 / world (DDP's average)
        ret[rank] = (n, eng.grads.clone(), local)
    finally:
        dist.destroy_process_group()


def test_gradient_all_reduce_world_size_2_gloo():
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29650 + os.getpid() % 200
    mp.spawn(_ddp_worker, args=(2, port, ret), nprocs=2, join=True)
    (n0, s0, l0), (n1, s1, l1) = ret[0], ret[1]
    assert n0 == n1 == 2
    assert torch.equal(s0, s1)                       # ranks stay in lock step
    assert torch.allclose(s0, l0 + l1)
    assert torch.allclose(s0 / n0, (l0 + l1) / 2)    # the average DDP would have produced


def test_single_process_exchange_is_the_identity():
    g = torch.arange(10.0)
    assert T.TrainEngine.all_reduce_grads(_Flat(g)) == 1 and torch.equal(g, torch.arange(10.0))
