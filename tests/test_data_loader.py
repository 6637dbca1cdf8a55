"""Batch collation for ``forward`` (reference data_loader.py:201-287): the product's
``DaftExprtDataCollate`` against the live reference class (build container) and against a plain restatement, errors, the
one-block staging layout; GPU: ``to_device`` feeds ``forward`` the same tensors as ``parse_batch`` on the padded batch."""
import pytest
import torch

import _cases as C
from daft_exprt_b200 import synthetic
from daft_exprt_b200.data_loader import DaftExprtDataCollate
from oracle import ref_harness as R


def _items(n, seed, with_emb=True):
    hp = C.hparams()
    items = []
    for i in range(n):
        u = synthetic.utterance(hp, i, seed=seed, lmin=3, lmax=40)
        items.append([u["symbols"].int(), u["durations_float"], u["durations_int"].int(), u["symbols_energy"], u["symbols_pitch"],
                      u["frames_energy"], u["frames_pitch"], u["mel"], int(u["speaker_id"]), f"/feat/spk{int(u['speaker_id'])}",
                      f"file{i:03d}", u["spk_emb"] if with_emb else None])
    return hp, items


def _restated(hp, batch):
    """data_loader.py:214-287 restated with plain loops (test-local checker)."""
    lengths = torch.LongTensor([len(x[0]) for x in batch])
    in_len, order = torch.sort(lengths, dim=0, descending=True)
    B, L, T = len(batch), int(in_len[0]), max(x[7].size(1) for x in batch)
    out = [torch.zeros(B, L, dtype=torch.long), torch.zeros(B, L), torch.zeros(B, L, dtype=torch.long), torch.zeros(B, L),
           torch.zeros(B, L), in_len, torch.zeros(B, T), torch.zeros(B, T), torch.zeros(B, hp.n_mel_channels, T),
           torch.zeros(B, dtype=torch.long), torch.zeros(B, dtype=torch.long), [], [], torch.zeros(B, batch[0][11].numel())]
    for i, src in enumerate(order.tolist()):
        x = batch[src]
        n, t = len(x[0]), x[7].size(1)
        for k in range(5):
            out[k][i, :n] = x[k]
        out[6][i, :t], out[7][i, :t], out[8][i, :, :t] = x[5], x[6], x[7]
        out[9][i], out[10][i] = t, x[8]
        out[11].append(x[9])
        out[12].append(x[10])
        out[13][i] = x[11]
    return tuple(out)


def _same(a, b):
    assert len(a) == len(b) == 14
    for k in range(14):
        if k in (11, 12):
            assert a[k] == b[k], k
        else:
            assert a[k].dtype == b[k].dtype and torch.equal(a[k], b[k]), k


@pytest.mark.parametrize("n,seed", [(1, 3), (7, 11), (16, 12)])
def test_collate_equals_restatement(n, seed):
    hp, items = _items(n, seed)
    _same(DaftExprtDataCollate(hp, pinned=False)(items), _restated(hp, items))


@pytest.mark.skipif(not R.reference_available(), reason="/root/reference not present")
@pytest.mark.parametrize("n,seed", [(5, 21), (12, 22)])
def test_collate_equals_reference(n, seed):
    R.import_reference_generate()
    import daft_exprt.data_loader as ref_dl
    hp, items = _items(n, seed)
    _same(DaftExprtDataCollate(hp, pinned=False)(items), ref_dl.DaftExprtDataCollate(hp)(items))


def test_collate_errors_and_layout():
    hp, items = _items(4, 5, with_emb=False)
    with pytest.raises(ValueError, match="required for every sample"):
        DaftExprtDataCollate(hp, pinned=False)(items)
    hp, items = _items(4, 5)
    items[2][11] = None
    with pytest.raises(ValueError, match="Mixed presence of speaker embeddings"):
        DaftExprtDataCollate(hp, pinned=False)(items)
    hp, items = _items(4, 5)
    b = DaftExprtDataCollate(hp, pinned=False)(items)
    spans = sorted((off, n) for off, n, _, _ in b.layout.values())
    assert all(off % 256 == 0 for off, _ in spans)
    assert all(a + n <= nxt for (a, n), (nxt, _) in zip(spans, spans[1:])) and spans[-1][0] + spans[-1][1] <= b.block.numel()
    assert all(t.untyped_storage().data_ptr() == b.block.untyped_storage().data_ptr() for k, t in enumerate(b) if k not in (11, 12))


@pytest.mark.gpu
def test_to_device_feeds_forward_like_parse_batch():
    from daft_exprt_b200.model import DaftExprt
    hp, items = _items(9, 31)
    model = DaftExprt(hp)
    model.load_state_dict(C.state_dict())
    model = model.cuda().eval().set_mode("fp32")
    batch = DaftExprtDataCollate(hp)(items)
    assert batch.block.is_pinned()
    inputs, _ = model.parse_batch(0, batch)
    fast = batch.to_device(model.device)
    for a, b in zip(inputs, fast):
        assert a.dtype == b.dtype and torch.equal(a, b)
    out_a, out_b = model(inputs), model(fast)
    assert torch.equal(out_a[3][0], out_b[3][0]) and torch.equal(out_a[4], out_b[4])


def test_collated_batch_pickles_and_survives_dataloader_workers():
    """the reference's prepare_data_loaders uses this collate with num_workers >= 1 (data_loader.py:290-330);
    the tuple subclass must cross the worker queue and DataLoader's pin_memory re-wrap."""
    import pickle
    from torch.utils.data import DataLoader, Dataset
    from daft_exprt_b200.data_loader import CollatedBatch
    hp, items = _items(6, 41)
    batch = DaftExprtDataCollate(hp, pinned=False)(items)
    clone = pickle.loads(pickle.dumps(batch))
    assert isinstance(clone, CollatedBatch) and clone.layout == batch.layout
    _same(clone, batch)

    class _DS(Dataset):
        def __len__(self):
            return len(items)

        def __getitem__(self, i):
            return items[i]

    loader = DataLoader(_DS(), batch_size=3, shuffle=False, num_workers=1, collate_fn=DaftExprtDataCollate(hp))
    got = list(loader)
    assert len(got) == 2
    _same(got[0], DaftExprtDataCollate(hp, pinned=False)(items[:3]))
    _same(got[1], DaftExprtDataCollate(hp, pinned=False)(items[3:]))
    # a batch rebuilt without its block (what DataLoader(pin_memory=True) does to sequences) still moves
    bare = CollatedBatch(tuple(batch))
    moved = bare.to_device("cpu")
    assert len(moved) == 12 and torch.equal(moved[8], batch[8])
