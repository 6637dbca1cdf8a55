"""Bulk mel generation, batch-sharded across GPUs (BASELINE.json config 3; the reference's single-GPU
``fine_tune.py:62-129`` loop).

Utterances are independent and the weights are replicated, so the path shards with **no data-path
collective**: every rank derives the same deterministic plan from the utterance lengths alone
(global length sort → batches of ``batch_size`` → batches dealt to ranks so that the quadratic
attention cost balances), runs ``DaftExprt.forward`` on its own batches and keeps / writes its own
results.  ``torch.distributed`` is only used for the optional final gather of small summaries and
for the barrier that brackets timing.

Host-side logic only (lists of ints, index tensors); the forward pass is the model's.
"""
import queue
import threading

import torch


# ----------------------------------------------------------------------------------------------
# planning
# ----------------------------------------------------------------------------------------------
def batch_cost(in_lens, out_lens):
    """Relative cost of one padded batch: linear (conv/FFN) + quadratic (attention) terms of the
    closed-form FLOP model, frame side dominating."""
    c = 0.0
    for l, t in zip(in_lens, out_lens):
        c += 8.0 * (1_703_936.0 * t + 512.0 * t * t) + 7_570_944.0 * t + 4.0 * (1_703_936.0 * l + 512.0 * l * l)
    return c


def plan_batches(in_lens, out_lens, batch_size):
    """Global sort by symbol count, descending (ties: by index), then consecutive chunks.  Inside a
    batch the order is the reference collate's (``data_loader.py:217``: descending input length).
    Returns a list of index lists."""
    order = sorted(range(len(in_lens)), key=lambda i: (-int(in_lens[i]), i))
    return [order[i:i + batch_size] for i in range(0, len(order), batch_size)]


def assign_batches(batches, in_lens, out_lens, world_size):
    """Greedy longest-processing-time assignment of batches to ranks (deterministic: ties go to the
    lowest rank).  Returns ``world_size`` lists of batch indices."""
    costs = [batch_cost([in_lens[i] for i in b], [out_lens[i] for i in b]) for b in batches]
    order = sorted(range(len(batches)), key=lambda k: (-costs[k], k))
    load = [0.0] * world_size
    mine = [[] for _ in range(world_size)]
    for k in order:
        r = min(range(world_size), key=lambda q: (load[q], q))
        mine[r].append(k)
        load[r] += costs[k]
    for r in range(world_size):
        mine[r].sort()
    return mine


def plan_shards(in_lens, out_lens, batch_size, world_size):
    """Full plan: ``(batches, per_rank_batch_ids)``.  Every rank computes the same plan."""
    batches = plan_batches(in_lens, out_lens, batch_size)
    return batches, assign_batches(batches, in_lens, out_lens, world_size)


# ----------------------------------------------------------------------------------------------
# collation (reference ``data_loader.py:201-287`` for tensors already in memory)
# ----------------------------------------------------------------------------------------------
def collate(utterances, idx):
    """``utterances``: dict of per-utterance lists of CPU tensors with the reference's per-item
    fields (symbols (L,), durations_float (L,), durations_int (L,), symbols_energy (L,),
    symbols_pitch (L,), frames_energy (T,), frames_pitch (T,), mel (80, T), speaker_id (),
    spk_emb (192,)).  ``idx``: utterance ids of one batch, already in descending symbol count.
    Returns the 12-tuple of ``DaftExprt.forward`` (right-zero-padded, pinned)."""
    B = len(idx)
    L = max(int(utterances["symbols"][i].shape[0]) for i in idx)
    T = max(int(utterances["mel"][i].shape[1]) for i in idx)
    M = int(utterances["mel"][idx[0]].shape[0])
    pin = torch.cuda.is_available()
    z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, pin_memory=pin)   # noqa: E731
    symbols, dint = z(B, L, dt=torch.long), z(B, L, dt=torch.long)
    dfl, sen, spi = z(B, L), z(B, L), z(B, L)
    fen, fpi, mel = z(B, T), z(B, T), z(B, M, T)
    in_len, out_len, spk = z(B, dt=torch.long), z(B, dt=torch.long), z(B, dt=torch.long)
    emb = z(B, int(utterances["spk_emb"][idx[0]].shape[0]))
    for b, i in enumerate(idx):
        l, t = int(utterances["symbols"][i].shape[0]), int(utterances["mel"][i].shape[1])
        symbols[b, :l] = utterances["symbols"][i]
        dfl[b, :l] = utterances["durations_float"][i]
        dint[b, :l] = utterances["durations_int"][i]
        sen[b, :l] = utterances["symbols_energy"][i]
        spi[b, :l] = utterances["symbols_pitch"][i]
        fen[b, :t] = utterances["frames_energy"][i]
        fpi[b, :t] = utterances["frames_pitch"][i]
        mel[b, :, :t] = utterances["mel"][i]
        in_len[b], out_len[b] = l, t
        spk[b] = utterances["speaker_id"][i]
        emb[b] = utterances["spk_emb"][i]
    return (symbols, dfl, dint, sen, spi, in_len, fen, fpi, mel, out_len, spk, emb)


# ----------------------------------------------------------------------------------------------
# staged collation: one pinned block per batch, filled with numpy on worker threads
# ----------------------------------------------------------------------------------------------
_ALIGN = 256
_FORWARD_ORDER = ("symbols", "durations_float", "durations_int", "symbols_energy", "symbols_pitch", "input_lengths",
                  "frames_energy", "frames_pitch", "mel_specs", "output_lengths", "speaker_ids", "spk_embs")


def block_layout(B, L, T, M, E, total_frames=None):
    """Byte layout of one batch inside a staging block (the plan of ``data_loader.DaftExprtDataCollate``):
    ``{name: (offset, nbytes, torch dtype, shape)}`` and the total size.  ``total_frames`` (= Σ T_b): the frame side
    is PACKED (``include/daft_b200.h`` DX_IO_PACKED_FRAMES) — only valid frames are staged and cross PCIe."""
    i8, f4 = torch.int64, torch.float32
    if total_frames is None:
        frame_side = [("frames_energy", f4, (B, T)), ("frames_pitch", f4, (B, T)), ("mel_specs", f4, (B, M, T))]
    else:
        frame_side = [("frames_energy", f4, (total_frames,)), ("frames_pitch", f4, (total_frames,)),
                      ("mel_specs", f4, (total_frames * M,))]
    plan = [("symbols", i8, (B, L)), ("durations_int", i8, (B, L)), ("input_lengths", i8, (B,)), ("output_lengths", i8, (B,)),
            ("speaker_ids", i8, (B,)), ("durations_float", f4, (B, L)), ("symbols_energy", f4, (B, L)),
            ("symbols_pitch", f4, (B, L))] + frame_side + [("spk_embs", f4, (B, E))]
    layout, off = {}, 0
    for name, dt, shape in plan:
        n = 1
        for d in shape:
            n *= d
        nbytes = n * (8 if dt == i8 else 4)
        layout[name] = (off, nbytes, dt, shape)
        off += (nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
    return layout, max(off, _ALIGN)


def fill_block(block, utterances, idx, packed=False):
    """Collate the utterances ``idx`` (descending symbol count) into the uint8 staging tensor ``block`` (reference
    ``data_loader.py:214-263``: right zero padding).  numpy slice copies — they release the GIL for the large mel rows,
    so several batches can be collated side by side — and only the pad tails are zeroed, not the whole block.
    ``packed``: the frame side is packed (one copy per utterance, no padding).  Returns the layout."""
    import numpy as np
    B = len(idx)
    lens = [int(utterances["symbols"][i].shape[0]) for i in idx]
    tlens = [int(utterances["mel"][i].shape[1]) for i in idx]
    L, T = max(lens), max(tlens)
    M = int(utterances["mel"][idx[0]].shape[0])
    E = int(utterances["spk_emb"][idx[0]].shape[0])
    layout, total = block_layout(B, L, T, M, E, sum(tlens) if packed else None)
    assert total <= block.numel(), (total, block.numel())
    raw = block.numpy()
    v = {}
    for name, (off, nbytes, dt, shape) in layout.items():
        v[name] = raw[off:off + nbytes].view(np.int64 if dt == torch.int64 else np.float32).reshape(shape)
    per_l = (("symbols", "symbols"), ("durations_float", "durations_float"), ("durations_int", "durations_int"),
             ("symbols_energy", "symbols_energy"), ("symbols_pitch", "symbols_pitch"))
    for b, i in enumerate(idx):
        l, t = lens[b], tlens[b]
        for dst, src in per_l:
            row = v[dst][b]
            row[:l] = utterances[src][i].numpy()
            row[l:] = 0
        if packed:
            f0 = sum(tlens[:b])
            v["frames_energy"][f0:f0 + t] = utterances["frames_energy"][i].numpy()
            v["frames_pitch"][f0:f0 + t] = utterances["frames_pitch"][i].numpy()
            v["mel_specs"][M * f0:M * (f0 + t)] = utterances["mel"][i].numpy().reshape(-1)
        else:
            for dst, src in (("frames_energy", "frames_energy"), ("frames_pitch", "frames_pitch")):
                row = v[dst][b]
                row[:t] = utterances[src][i].numpy()
                row[t:] = 0
            mel = v["mel_specs"][b]
            mel[:, :t] = utterances["mel"][i].numpy()
            mel[:, t:] = 0
        v["input_lengths"][b], v["output_lengths"][b] = l, t
        v["speaker_ids"][b] = int(utterances["speaker_id"][i])
        v["spk_embs"][b] = utterances["spk_emb"][i].numpy()
    return layout


class PointerTable:
    """Host pointers and lengths of every utterance of an in-memory stream, gathered once, so that a batch is collated
    by ONE native call (``dx_host_collate``: memcpy / memset outside the GIL, several host threads) instead of ~700
    Python-level slice assignments.  ``utterances``: the dict of per-utterance lists ``collate`` takes; tensors must be
    contiguous CPU tensors of the reference's dtypes (int64 / float32) and are kept alive by this object."""

    _SRC = (("symbols", torch.int64), ("durations_float", torch.float32), ("durations_int", torch.int64),
            ("symbols_energy", torch.float32), ("symbols_pitch", torch.float32), ("frames_energy", torch.float32),
            ("frames_pitch", torch.float32), ("mel", torch.float32), ("speaker_id", torch.int64), ("spk_emb", torch.float32))

    def __init__(self, utterances, n=None):
        import numpy as np
        n = len(utterances["symbols"]) if n is None else n
        self.n = n
        self.keep = []
        self.ptrs = np.zeros((n, len(self._SRC)), dtype=np.uint64)
        self.lens = np.zeros((n, 2), dtype=np.int32)
        for i in range(n):
            for f, (name, dt) in enumerate(self._SRC):
                t = utterances[name][i]
                if t.dtype != dt or not t.is_contiguous() or t.device.type != "cpu":
                    t = t.to(device="cpu", dtype=dt).contiguous()
                self.keep.append(t)
                self.ptrs[i, f] = t.data_ptr()
            self.lens[i, 0] = utterances["symbols"][i].shape[0]
            self.lens[i, 1] = utterances["mel"][i].shape[1]
        self.M = int(utterances["mel"][0].shape[0])
        self.E = int(utterances["spk_emb"][0].numel())

    def fill(self, block, idx, threads=1, packed=False):
        """Native collate of the utterances ``idx`` into the uint8 staging tensor ``block``; returns the layout.
        ``packed``: frame side in the DX_IO_PACKED_FRAMES layout (``dx_host_collate_packed``)."""
        import ctypes
        import numpy as np
        from . import _lib
        ids = np.asarray(idx, dtype=np.int64)
        lens = np.ascontiguousarray(self.lens[ids])
        ptrs = np.ascontiguousarray(self.ptrs[ids])
        B, L, T = len(idx), int(lens[:, 0].max()), int(lens[:, 1].max())
        layout, total = block_layout(B, L, T, self.M, self.E, int(lens[:, 1].sum()) if packed else None)
        assert total <= block.numel(), (total, block.numel())
        offs = np.array([layout[name][0] for name in _FORWARD_ORDER], dtype=np.int64)
        lib = _lib.load()
        fn = lib.dx_host_collate_packed if packed else lib.dx_host_collate
        _lib.check(fn(B, L, T, self.M, self.E, ctypes.c_void_p(ptrs.ctypes.data), ctypes.c_void_p(lens.ctypes.data),
                      ctypes.c_void_p(block.data_ptr()), ctypes.c_void_p(offs.ctypes.data), int(threads)),
                   "dx_host_collate_packed" if packed else "dx_host_collate")
        return layout


def block_views(buf, layout):
    """The 12 tensors of ``DaftExprt.forward`` as views of a staging block or of its device mirror."""
    out = []
    for name in _FORWARD_ORDER:
        off, nbytes, dt, shape = layout[name]
        out.append(buf[off:off + nbytes].view(dt).view(shape))
    return tuple(out)


# ----------------------------------------------------------------------------------------------
# execution
# ----------------------------------------------------------------------------------------------
class BulkRunner:
    """Runs ``forward`` over this rank's batches as a pipeline (``fine_tune.py:62-129`` without its blocking steps):

    * ``workers`` threads collate the next batches into a ring of pinned staging blocks — natively
      (``PointerTable.fill`` → ``dx_host_collate``, no GIL, ``collate_threads`` host threads per batch) when the
      utterances are in memory, else ``fill_block``,
    * the main thread sends a block H2D as ONE copy on a side stream and runs ``forward`` on views of its device mirror,
    * batch k-1's mel goes D2H into pinned memory and a writer thread hands ``sink(utt_id, mel (80, T_b) numpy view)``
      the per-item crops (the reference's ``fine_tune.py:77-80`` without its blocking ``.cpu()``).

    On a CPU model (tests) the same plan runs unstaged."""

    def __init__(self, model, sink=None, device=None, workers=2, slots=6, collate_threads=4, packed=True):
        self.model = model
        self.sink = sink
        self.packed = bool(packed)      # only valid frames are staged, copied H2D and copied back (DX_IO_PACKED_*)
        self.device = device if device is not None else model.device
        self.frames = 0
        self._writer_error = None
        self.workers, self.slots, self.collate_threads = max(1, int(workers)), max(3, int(slots)), max(1, int(collate_threads))

    def run(self, utterances, batches, my_batch_ids):
        if self.device.type != "cuda":
            return self._run_unstaged(utterances, batches, my_batch_ids)
        from concurrent.futures import ThreadPoolExecutor
        dev = self.device
        if not my_batch_ids:
            return self.frames
        # in-memory streams (dict of lists / PointerTable) are collated natively; lazy streams fall back to numpy copies
        table = utterances if isinstance(utterances, PointerTable) else None
        if table is None and isinstance(utterances, dict) and isinstance(utterances.get("symbols"), (list, tuple)):
            table = PointerTable(utterances)
        # capacity of a staging block: the largest batch of this shard
        cap = 0
        for k in my_batch_ids:
            idx = batches[k]
            if table is not None:
                L, T = int(table.lens[idx, 0].max()), int(table.lens[idx, 1].max())
                M, E = table.M, table.E
            else:
                L = max(int(utterances["symbols"][i].shape[0]) for i in idx)
                T = max(int(utterances["mel"][i].shape[1]) for i in idx)
                M, E = int(utterances["mel"][idx[0]].shape[0]), int(utterances["spk_emb"][idx[0]].shape[0])
            mel_channels = M
            cap = max(cap, block_layout(len(idx), L, T, M, E)[1])      # padded layout: an upper bound for the packed one
        S = self.slots
        blocks = [torch.empty(cap, dtype=torch.uint8, pin_memory=True) for _ in range(S)]
        mirrors = [torch.empty(cap, dtype=torch.uint8, device=dev) for _ in range(S)]
        h2d_done = [None] * S                 # the block may be refilled once its last H2D has completed
        fwd_done = [None] * S                 # the mirror may be overwritten once the forward reading it has completed
        copy_stream = torch.cuda.Stream(dev)
        d2h_stream = torch.cuda.Stream(dev)
        q = queue.Queue(maxsize=4)
        writer = threading.Thread(target=self._writer, args=(q,), daemon=True)
        writer.start()

        packed = self.packed

        def collate_into(pos):
            s = pos % S
            if h2d_done[s] is not None:
                h2d_done[s].synchronize()
            if table is not None:
                return table.fill(blocks[s], batches[my_batch_ids[pos]], self.collate_threads, packed)
            return fill_block(blocks[s], utterances, batches[my_batch_ids[pos]], packed)

        n = len(my_batch_ids)
        depth = min(S - 1, self.workers + 1)
        try:
            with ThreadPoolExecutor(max_workers=self.workers) as pool:
                futs = {pos: pool.submit(collate_into, pos) for pos in range(min(depth, n))}
                for pos in range(n):
                    if self._writer_error is not None:       # the sink failed: stop producing, re-raised below
                        break
                    s = pos % S
                    layout = futs.pop(pos).result()
                    nbytes = max(o + b for o, b, _, _ in layout.values())
                    with torch.cuda.stream(copy_stream):
                        if fwd_done[s] is not None:
                            copy_stream.wait_event(fwd_done[s])
                        mirrors[s][:nbytes].copy_(blocks[s][:nbytes], non_blocking=True)
                        h2d_done[s] = torch.cuda.Event()
                        h2d_done[s].record(copy_stream)
                    if pos + depth < n:       # its slot is the one whose H2D was enqueued an iteration ago at the latest
                        futs[pos + depth] = pool.submit(collate_into, pos + depth)
                    torch.cuda.current_stream(dev).wait_event(h2d_done[s])
                    out_len = block_views(blocks[s], layout)[9].clone()
                    views = block_views(mirrors[s], layout)
                    if packed:
                        T_max, total = int(out_len.max()), int(out_len.sum())
                        out = self.model(views, packed_io=(T_max, total))
                    else:
                        out = self.model(views)
                    fwd_done[s] = torch.cuda.Event()
                    fwd_done[s].record()
                    mel = out[3][0]
                    hmel = torch.empty(mel.shape, dtype=mel.dtype, pin_memory=True)
                    # the result leaves on its own stream: on the compute stream the next batch's forward would queue behind
                    # this copy (8.7 MB per cfg2 batch: 0.2-0.35 ms of a 1.25 ms step)
                    with torch.cuda.stream(d2h_stream):
                        d2h_stream.wait_event(fwd_done[s])
                        hmel.copy_(mel, non_blocking=True)
                        done = torch.cuda.Event()
                        done.record(d2h_stream)
                    mel.record_stream(d2h_stream)       # the allocator must not hand the block out again before the copy ran
                    self.frames += int(out_len.sum())
                    # packed output: the writer thread cuts the per-utterance (80, T_b) views out of the pinned buffer
                    q.put((batches[my_batch_ids[pos]], out_len, hmel, done, mel_channels if packed else 0))
                for f in futs.values():       # a failed run must not leave collate jobs writing into freed blocks
                    f.cancel()
        finally:
            self._finish_writer(q, writer)
        return self.frames

    def _run_unstaged(self, utterances, batches, my_batch_ids):
        q = queue.Queue(maxsize=4)
        writer = threading.Thread(target=self._writer, args=(q,), daemon=True)
        writer.start()
        try:
            for k in my_batch_ids:
                if self._writer_error is not None:
                    break
                host = collate(utterances, batches[k])
                out = self.model(host)
                self.frames += int(host[9].sum())
                q.put((batches[k], host[9], out[3][0], None, 0))
        finally:
            self._finish_writer(q, writer)
        return self.frames

    def _writer(self, q):
        """Writer thread: hands the per-utterance crops to ``sink``.  An exception from ``sink`` (disk full, bad path) is
        kept in ``self._writer_error``; the thread keeps DRAINING the queue (without calling ``sink`` again) so that the
        producer never blocks on a full queue, and ``run`` re-raises the error after ``join``."""
        while True:
            item = q.get()
            if item is None:
                return
            if self._writer_error is not None:
                continue
            try:
                ids, out_len, hmel, done, packed_channels = item
                if done is not None:
                    done.synchronize()
                if self.sink is not None:
                    if packed_channels:          # packed output: per-utterance (80, T_b) numpy views, the crop is free
                        arr, o = hmel.numpy(), 0
                        for i, n in zip(ids, out_len.tolist()):
                            self.sink(i, arr[o:o + packed_channels * n].reshape(packed_channels, n))
                            o += packed_channels * n
                    else:
                        arr = hmel.numpy()
                        for b, i in enumerate(ids):
                            self.sink(i, arr[b, :, :int(out_len[b])])
            except BaseException as e:          # noqa: BLE001 - re-raised in run()
                self._writer_error = e

    def _finish_writer(self, q, writer):
        q.put(None)
        writer.join()
        if self._writer_error is not None:
            err, self._writer_error = self._writer_error, None
            raise err
