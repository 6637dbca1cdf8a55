"""Host logic of the multi-GPU path: the shard plan and the sharded run, world_size 2 over
``gloo`` on CPU.  The "model" here is the oracle port (tests may use it as the checker) or a cheap stand-in."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _cases as C
from daft_exprt_b200 import bulk, synthetic
from oracle import daft_oracle as O


def test_plan_is_a_partition_and_balanced():
    hp = C.hparams()
    in_lens, out_lens = synthetic.stream_lengths(hp, 1000, seed=43)
    for world in (1, 2, 4, 8):
        batches, mine = bulk.plan_shards(in_lens, out_lens, 64, world)
        seen = sorted(i for b in batches for i in b)
        assert seen == list(range(1000))                       # every utterance exactly once
        assert sorted(k for m in mine for k in m) == list(range(len(batches)))
        for b in batches:                                      # collate order: descending symbol count
            assert [in_lens[i] for i in b] == sorted((in_lens[i] for i in b), reverse=True)
        loads = [sum(bulk.batch_cost([in_lens[i] for i in batches[k]], [out_lens[i] for i in batches[k]]) for k in m)
                 for m in mine]
        assert max(loads) <= 1.25 * (sum(loads) / world) + 1e-9
    # deterministic
    assert bulk.plan_shards(in_lens, out_lens, 64, 4) == bulk.plan_shards(in_lens, out_lens, 64, 4)


def test_collate_matches_forward_batch_contract():
    hp = C.hparams()
    stream = synthetic.UtteranceStream(hp, 10, seed=7, lmin=3, lmax=12)
    in_lens, out_lens = stream.lengths()
    batches = bulk.plan_batches(in_lens, out_lens, 4)
    batch = bulk.collate(stream, batches[0])
    assert len(batch) == 12
    assert batch[5].tolist() == sorted(batch[5].tolist(), reverse=True)
    assert torch.equal(batch[2].sum(dim=1), batch[9])           # data_loader.py:165-168 invariant
    assert batch[8].shape == (4, 80, int(batch[9].max()))
    for b in range(4):                                          # right-zero-padded
        assert batch[0][b, int(batch[5][b]):].abs().sum() == 0
        assert batch[8][b, :, int(batch[9][b]):].abs().sum() == 0


class _OracleModel:
    """Stand-in with the model's call signature, backed by the oracle port (CPU)."""

    def __init__(self, hp, sd):
        self.hp, self.sd = hp, sd
        self.device = torch.device("cpu")

    def __call__(self, batch):
        return O.forward(self.sd, self.hp, tuple(t.clone() for t in batch))


def _worker(rank, world, port, n_utt, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    hp = C.hparams()
    stream = synthetic.UtteranceStream(hp, n_utt, seed=11, lmin=3, lmax=10, dmax=3)
    in_lens, out_lens = stream.lengths()
    batches, mine = bulk.plan_shards(in_lens, out_lens, 3, world)
    got = {}
    runner = bulk.BulkRunner(_OracleModel(hp, C.state_dict()), sink=lambda i, mel: got.__setitem__(i, torch.from_numpy(mel.copy())))
    frames = runner.run(stream, batches, mine[rank])
    # no data-path collective: only the summary is gathered
    tot = torch.tensor([frames], dtype=torch.long)
    dist.all_reduce(tot)
    gathered = [None] * world
    dist.all_gather_object(gathered, sorted(got))
    if rank == 0:
        ret["frames"] = int(tot)
        ret["ids"] = gathered
    ret[f"mels{rank}"] = {i: m for i, m in got.items()}
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_sharded_run_equals_single_process():
    n_utt, world = 10, 2
    port = 29500 + (os.getpid() % 2000)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, n_utt, ret), nprocs=world, join=True)
    hp = C.hparams()
    stream = synthetic.UtteranceStream(hp, n_utt, seed=11, lmin=3, lmax=10, dmax=3)
    in_lens, out_lens = stream.lengths()
    assert ret["frames"] == sum(out_lens)
    ids = ret["ids"]
    assert sorted(ids[0] + ids[1]) == list(range(n_utt)) and not set(ids[0]) & set(ids[1])
    # same batches on one process → identical mels (batch composition is part of the contract)
    batches, _ = bulk.plan_shards(in_lens, out_lens, 3, 1)
    single = {}
    bulk.BulkRunner(_OracleModel(hp, C.state_dict()), sink=lambda i, mel: single.__setitem__(i, torch.from_numpy(mel.copy()))).run(
        stream, batches, list(range(len(batches))))
    merged = {}
    merged.update(ret["mels0"]); merged.update(ret["mels1"])
    for i in range(n_utt):
        assert merged[i].shape == (80, out_lens[i])
        # worker processes run the oracle with 2 intra-op threads, this process with the default: last-bit only
        assert torch.allclose(merged[i], single[i], atol=1e-5, rtol=0)


def test_staged_block_collate_equals_plain_collate():
    """The worker-thread collate (numpy copies into a reused staging block, pad tails zeroed) yields the tensors of
    the plain collate, also when the block held a larger batch before."""
    hp = C.hparams()
    stream = synthetic.UtteranceStream(hp, 12, seed=7, lmin=3, lmax=14)
    in_lens, out_lens = stream.lengths()
    batches = bulk.plan_batches(in_lens, out_lens, 4)
    block = torch.full((1 << 20,), 255, dtype=torch.uint8)          # dirty: nothing may leak into the pad regions
    for idx in batches:
        layout = bulk.fill_block(block, stream, idx)
        got = bulk.block_views(block, layout)
        ref = bulk.collate(stream, idx)
        assert len(got) == len(ref) == 12
        for g, r in zip(got, ref):
            assert g.dtype == r.dtype and g.shape == r.shape and torch.equal(g, r)


def test_native_collate_equals_plain_collate():
    """dx_host_collate (memcpy / memset in the library, 1 and 3 host threads) against the plain collate, on a dirty
    block; bad lengths are refused."""
    import pytest as _pt
    from daft_exprt_b200 import _lib
    hp = C.hparams()
    lazy = synthetic.UtteranceStream(hp, 11, seed=5, lmin=2, lmax=15)
    stream = {f: [lazy[f][i] for i in range(11)] for f in synthetic.UtteranceStream.FIELDS}
    in_lens, out_lens = lazy.lengths()
    table = bulk.PointerTable(stream)
    for threads in (1, 3):
        block = torch.full((1 << 20,), 170, dtype=torch.uint8)
        for idx in bulk.plan_batches(in_lens, out_lens, 4):
            layout = table.fill(block, idx, threads)
            for g, r in zip(bulk.block_views(block, layout), bulk.collate(stream, idx)):
                assert g.dtype == r.dtype and g.shape == r.shape and torch.equal(g, r)
    table.lens[3, 1] = -1
    with _pt.raises(_lib.DxError):
        table.fill(torch.zeros(1 << 20, dtype=torch.uint8), [3, 4], 1)


def test_packed_collate_equals_packed_padded_batch():
    """dx_host_collate_packed / fill_block(packed=True): the frame side holds exactly the valid frames, per-utterance
    (80, T_b) blocks, and equals layouts.pack_frames of the padded collate bit for bit; the symbol side is unchanged."""
    from daft_exprt_b200 import layouts
    hp = C.hparams()
    lazy = synthetic.UtteranceStream(hp, 13, seed=6, lmin=2, lmax=15)
    stream = {f: [lazy[f][i] for i in range(13)] for f in synthetic.UtteranceStream.FIELDS}
    in_lens, out_lens = lazy.lengths()
    table = bulk.PointerTable(stream)
    for idx in bulk.plan_batches(in_lens, out_lens, 5):
        ref = bulk.collate(stream, idx)
        e, p, m = layouts.pack_frames(ref[6], ref[7], ref[8], ref[9])
        for fill in (lambda blk: table.fill(blk, idx, 2, packed=True), lambda blk: bulk.fill_block(blk, stream, idx, packed=True)):
            block = torch.full((1 << 20,), 85, dtype=torch.uint8)
            got = bulk.block_views(block, fill(block))
            for k in (0, 1, 2, 3, 4, 5, 9, 10, 11):
                assert torch.equal(got[k], ref[k]), k
            assert torch.equal(got[6], e) and torch.equal(got[7], p) and torch.equal(got[8], m)
        # round trip of the helpers
        assert torch.equal(layouts.unpack_mel(m, ref[9], ref[8].shape[2]), ref[8])
        views = layouts.mel_views(m, ref[9])
        for b, v in enumerate(views):
            assert torch.equal(v, ref[8][b, :, :int(ref[9][b])])


def test_failing_sink_is_reported_not_hung():
    """an exception in sink() used to kill the writer thread silently and hang run() on the bounded queue."""
    hp = C.hparams()
    stream = synthetic.UtteranceStream(hp, 40, seed=8, lmin=2, lmax=6)
    in_lens, out_lens = stream.lengths()
    batches, mine = bulk.plan_shards(in_lens, out_lens, 2, 1)      # 20 batches > queue depth 4

    class _Cheap:
        device = torch.device("cpu")

        def __call__(self, batch):
            B, T = batch[8].shape[0], batch[8].shape[2]
            return None, None, None, [torch.zeros(B, 80, T), batch[9]], None

    calls = []

    def sink(i, mel):
        calls.append(i)
        if len(calls) == 3:
            raise OSError("disk full")

    runner = bulk.BulkRunner(_Cheap(), sink=sink)
    with pytest.raises(OSError, match="disk full"):
        runner.run(stream, batches, mine[0])
    assert len(calls) == 3                       # the sink is not called again after it failed
    # the runner is reusable afterwards
    calls.clear()
    runner2 = bulk.BulkRunner(_Cheap(), sink=lambda i, mel: None)
    assert runner2.run(stream, batches, mine[0]) == sum(out_lens)
