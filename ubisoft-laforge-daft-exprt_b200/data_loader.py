"""Batch collation for ``DaftExprt.forward`` (the loader side of ``fine_tune``).

Mirrors the reference's ``DaftExprtDataCollate`` (``src/daft_exprt/data_loader.py:201-287``): same constructor, same call
signature, same 14-tuple, same errors.  Input: a list of per-item lists as ``DaftExprtDataLoader.__getitem__`` returns them
(``data_loader.py:123-178``)::

    [symbols (L,) int, durations_float (L,), durations_int (L,) int, symbols_energy (L,), symbols_pitch (L,),
     frames_energy (T,), frames_pitch (T,), mel_spec (n_mel, T), speaker_id, features_dir, feature_file, spk_emb (E,) | None]

What differs is where the bytes land: the reference allocates 11 pageable tensors and fills them row by row
(``:214-259``); here every padded tensor of the batch is a typed view of ONE staging block (pinned when CUDA is there), so
the whole batch crosses PCIe as a single asynchronous copy (``to_device``), and the device tensors handed to ``forward``
are views of its device mirror.  The dataset / file side of the loader (``data_loader.py:14-199,290-345``: text files,
``.npy`` features, DistributedSampler) is host I/O and not part of this row; ``bulk.plan_shards`` is the sharded,
length-bucketed sampler.
"""
import numpy as np
import torch
import torch.utils.data

_ALIGN = 256


class CollatedBatch(tuple):
    """The reference's 14-tuple, plus the staging block it is a view of (``.block``, ``.layout``)."""

    def __new__(cls, items, block=None, layout=None):
        self = super().__new__(cls, items)
        self.block, self.layout = block, layout
        return self

    def __reduce__(self):
        # picklable (DataLoader workers send batches through a queue); the items are re-sent as plain tensors and the
        # staging block with them, so a batch that crossed a process boundary still moves as one copy
        return (CollatedBatch, (tuple(self), self.block, self.layout))

    def to_device(self, device, non_blocking=True):
        """One H2D copy of the staging block; returns the 12 tensors of ``DaftExprt.forward`` (``parse_batch`` order:
        the reference's batch indices 0-10 and 13) as views of the device mirror.  A batch rebuilt without its block
        (``DataLoader(pin_memory=True)`` re-wraps sequences as ``type(batch)(items)``) is moved tensor by tensor."""
        if self.block is None or not self._views_alias_block():
            return tuple(self[i].to(device, non_blocking=non_blocking) for i in (0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 13))
        mirror = torch.empty(self.block.numel(), dtype=torch.uint8, device=device)
        mirror.copy_(self.block, non_blocking=non_blocking)
        out = []
        for name in ("symbols", "durations_float", "durations_int", "symbols_energy", "symbols_pitch", "input_lengths",
                     "frames_energy", "frames_pitch", "mel_specs", "output_lengths", "speaker_ids", "spk_embs"):
            off, nbytes, dt, shape = self.layout[name]
            out.append(mirror[off:off + nbytes].view(dt).view(shape))
        return tuple(out)


def _views_alias(block, t):
    lo, hi = block.data_ptr(), block.data_ptr() + block.numel()
    return lo <= t.data_ptr() < hi


CollatedBatch._views_alias_block = lambda self: _views_alias(self.block, self[8])   # mel_specs: the bulk of the bytes


class DaftExprtDataCollate():
    ''' Zero-pads model inputs and targets (reference data_loader.py:201-287) '''

    def __init__(self, hparams, pinned=None):
        self.hparams = hparams
        self.pinned = pinned          # None: pinned when CUDA is there and this is not a (forked) DataLoader worker

    def _use_pinned(self):
        if self.pinned is not None:
            return bool(self.pinned)
        # pinning initialises CUDA, which a forked DataLoader worker cannot do ("cannot re-initialize CUDA in forked
        # subprocess"): workers fill pageable blocks, the main process pins them (DataLoader(pin_memory=True)) or not
        if torch.utils.data.get_worker_info() is not None:
            return False
        return torch.cuda.is_available()

    def __call__(self, batch):
        B = len(batch)
        # the reference's own sort call (data_loader.py:216-217), so ties come out in the same order
        input_lengths, ids_sorted_decreasing = torch.sort(torch.LongTensor([len(x[0]) for x in batch]), dim=0, descending=True)
        order = ids_sorted_decreasing.tolist()
        L = int(input_lengths[0])
        T = max([int(x[7].shape[1]) for x in batch])
        M = self.hparams.n_mel_channels

        spk_embs = [batch[idx][11] for idx in order]
        if all(item is None for item in spk_embs):
            raise ValueError(
                "Speaker embeddings (spk_embs) required for every sample. "
                "Run training.py pre_process to compute ECAPA embeddings and ensure .spk_emb.npy files exist in the feature directories."
            )
        if any(item is None for item in spk_embs):
            raise ValueError("Mixed presence of speaker embeddings in batch. All samples must have .spk_emb.npy.")
        E = int(np.prod(tuple(spk_embs[0].shape)))

        i8, f4 = torch.int64, torch.float32
        plan = [("symbols", i8, (B, L)), ("durations_int", i8, (B, L)), ("input_lengths", i8, (B,)), ("output_lengths", i8, (B,)),
                ("speaker_ids", i8, (B,)), ("durations_float", f4, (B, L)), ("symbols_energy", f4, (B, L)),
                ("symbols_pitch", f4, (B, L)), ("frames_energy", f4, (B, T)), ("frames_pitch", f4, (B, T)),
                ("mel_specs", f4, (B, M, T)), ("spk_embs", f4, (B, E))]
        layout, off = {}, 0
        for name, dt, shape in plan:
            nbytes = int(np.prod(shape)) * (8 if dt == i8 else 4)
            layout[name] = (off, nbytes, dt, shape)
            off += (nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
        block = torch.zeros(max(off, _ALIGN), dtype=torch.uint8, pin_memory=self._use_pinned())
        t = {name: block[o:o + n].view(dt).view(shape) for name, (o, n, dt, shape) in layout.items()}

        t["input_lengths"].copy_(input_lengths)
        feature_dirs, feature_files = [], []
        for i, src in enumerate(order):
            item = batch[src]
            n = item[0].shape[0]
            t["symbols"][i, :n] = torch.as_tensor(item[0])
            t["durations_float"][i, :item[1].shape[0]] = torch.as_tensor(item[1])
            t["durations_int"][i, :item[2].shape[0]] = torch.as_tensor(item[2])
            t["symbols_energy"][i, :item[3].shape[0]] = torch.as_tensor(item[3])
            t["symbols_pitch"][i, :item[4].shape[0]] = torch.as_tensor(item[4])
            t["speaker_ids"][i] = int(item[8])
            mel = torch.as_tensor(item[7])
            t["frames_energy"][i, :item[5].shape[0]] = torch.as_tensor(item[5])
            t["frames_pitch"][i, :item[6].shape[0]] = torch.as_tensor(item[6])
            t["mel_specs"][i, :, :mel.shape[1]] = mel
            t["output_lengths"][i] = mel.shape[1]
            feature_dirs.append(item[9])
            feature_files.append(item[10])
            t["spk_embs"][i] = torch.as_tensor(spk_embs[i]).reshape(-1)

        return CollatedBatch(
            (t["symbols"], t["durations_float"], t["durations_int"], t["symbols_energy"], t["symbols_pitch"], t["input_lengths"],
             t["frames_energy"], t["frames_pitch"], t["mel_specs"], t["output_lengths"], t["speaker_ids"], feature_dirs,
             feature_files, t["spk_embs"]), block, layout)
