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
# execution
# ----------------------------------------------------------------------------------------------
class BulkRunner:
    """Runs ``forward`` over this rank's batches with a two-deep pipeline: while batch k computes,
    batch k+1 is collated + copied H2D on a side stream and batch k-1's mel goes D2H into pinned
    memory and is handed to ``sink(utt_id, mel (80, T_b) numpy view)`` on a writer thread — the
    per-item crop of the reference (``fine_tune.py:77-80``) without its blocking ``.cpu()``."""

    def __init__(self, model, sink=None, device=None):
        self.model = model
        self.sink = sink
        self.device = device if device is not None else model.device
        self.frames = 0

    def run(self, utterances, batches, my_batch_ids):
        dev = self.device
        use_cuda = dev.type == "cuda"
        copy_stream = torch.cuda.Stream(dev) if use_cuda else None
        q = queue.Queue(maxsize=4)
        writer = threading.Thread(target=self._writer, args=(q,), daemon=True)
        writer.start()

        def stage(k):
            host = collate(utterances, batches[k])
            if not use_cuda:
                return host, host, None
            with torch.cuda.stream(copy_stream):
                d = tuple(t.to(dev, non_blocking=True) for t in host)
                ev = torch.cuda.Event()
                ev.record(copy_stream)
            return host, d, ev

        nxt = stage(my_batch_ids[0]) if my_batch_ids else None
        for pos, k in enumerate(my_batch_ids):
            host, d, ev = nxt
            if ev is not None:
                torch.cuda.current_stream(dev).wait_event(ev)
            out = self.model(d)
            nxt = stage(my_batch_ids[pos + 1]) if pos + 1 < len(my_batch_ids) else None
            mel = out[3][0]
            if use_cuda:
                hmel = torch.empty(mel.shape, dtype=mel.dtype, pin_memory=True)
                hmel.copy_(mel, non_blocking=True)
                done = torch.cuda.Event()
                done.record()
            else:
                hmel, done = mel, None
            self.frames += int(host[9].sum())
            q.put((batches[k], host[9], hmel, done))
        q.put(None)
        writer.join()
        return self.frames

    def _writer(self, q):
        while True:
            item = q.get()
            if item is None:
                return
            ids, out_len, hmel, done = item
            if done is not None:
                done.synchronize()
            if self.sink is not None:
                arr = hmel.numpy()
                for b, i in enumerate(ids):
                    self.sink(i, arr[b, :, :int(out_len[b])])
