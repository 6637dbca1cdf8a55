"""Measured numbers for the BASELINE.json configs that are not the bench headline (cfg1 latency, cfg3 bulk subset
through BulkRunner, cfg4 long form), acoustic model and vocoder, on one B200.  One JSON object on stdout.

    python tools/bench_configs.py > gpurun_out/configs.json
"""
import json
import os
import statistics
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from daft_exprt_b200 import bulk, synthetic, vocoder as VC
from daft_exprt_b200.hparams import default_hparams
from daft_exprt_b200.model import DaftExprt
from daft_exprt_b200.weights import make_state_dict

dev = torch.device("cuda", 0)
hp = default_hparams()
model = DaftExprt(hp, mode="bf16").to(dev)
model.load_state_dict(make_state_dict(hp, 1234))
model.eval()
gen = VC.HiFiGANGenerator(VC.DEFAULT_CONFIG)
gen.load_state_dict(VC.synthetic_state_dict(seed=4321))
gen.to(dev).eval().set_mode("bf16")
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timed(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return statistics.median(ms)


out = {}
with torch.no_grad():
    for name in ("cfg1", "cfg4_b1", "cfg4_b8"):
        batch = tuple(t.to(dev) for t in synthetic.config_batch(hp, name))
        frames = int(batch[9].sum())
        ms = timed(lambda: model(batch), 20)
        mel = model(batch)[3][0].contiguous()
        ms_v = timed(lambda: gen(mel, batch[9], total_frames=frames), 10)
        out[name] = {"B": int(batch[0].shape[0]), "L": int(batch[0].shape[1]), "T": int(mel.shape[2]), "valid_frames": frames,
                     "forward_ms": ms, "forward_frames_per_s": frames / ms * 1e3,
                     "vocoder_ms": ms_v, "vocoder_samples_per_s": frames * 256 / ms_v * 1e3,
                     "audio_seconds_per_call": frames * 256 / 22050.0}
    # p50 wall-clock latency of one cfg1 utterance: mel (forward) and mel + waveform incl. the D2H of the audio
    batch = tuple(t.to(dev) for t in synthetic.config_batch(hp, "cfg1"))
    lat = []
    for i in range(300):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        mel = model(batch)[3][0]
        audio = gen(mel, batch[9], total_frames=300).cpu()
        lat.append((time.perf_counter() - t0) * 1e3)
    out["cfg1"]["mel_plus_wave_latency_ms_p50"] = statistics.median(lat[50:])
    # cfg3: a bounded subset of the 100k-utterance stream through the pipelined bulk runner (collate + H2D + forward +
    # D2H + per-utterance crop), one GPU
    n = 4096
    lazy = synthetic.UtteranceStream(hp, n, seed=43)
    in_lens, out_lens = lazy.lengths()
    # utterances already in memory (what a feature cache / DataLoader workers deliver): the runner, not the synthetic
    # generator, is what is timed
    stream = {f: [None] * n for f in synthetic.UtteranceStream.FIELDS}
    for i in range(n):
        u = synthetic.utterance(hp, i, 43)
        for f in stream:
            stream[f][i] = u[f]
    batches, mine = bulk.plan_shards(in_lens, out_lens, 64, 1)
    sunk = [0]

    def sink(i, mel):
        sunk[0] += mel.shape[1]
    res = {}
    table = bulk.PointerTable(stream)
    for workers, cthreads in ((1, 1), (1, 4), (2, 4), (2, 8)):
        runner = bulk.BulkRunner(model, sink=sink, workers=workers, slots=6, collate_threads=cthreads)
        runner.run(table, batches, mine[0][:8])          # warm-up
        runner.frames, sunk[0] = 0, 0
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        frames = runner.run(table, batches, mine[0])
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        assert sunk[0] == frames == sum(out_lens)
        res[f"workers_{workers}_collate_threads_{cthreads}"] = {"seconds": dt, "frames_per_s": frames / dt, "utterances_per_s": n / dt}
    # the plain single-thread collate of round 1 (12 pinned tensors per batch, torch slice assignments), same pipeline depth
    t0 = time.perf_counter()
    for k in mine[0][:16]:
        bulk.collate(stream, batches[k])
    t_plain = (time.perf_counter() - t0) / 16
    blk = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
    t0 = time.perf_counter()
    for k in mine[0][:16]:
        bulk.fill_block(blk, stream, batches[k])
    t_block = (time.perf_counter() - t0) / 16
    t_nat = {}
    for th in (1, 4, 8):
        t0 = time.perf_counter()
        for k in mine[0][:16]:
            table.fill(blk, batches[k], th)
        t_nat[f"native_{th}_threads"] = (time.perf_counter() - t0) / 16 * 1e3
    out["cfg3_subset"] = {"utterances": n, "batches": len(mine[0]), "valid_frames": int(sum(out_lens)), "host_cores": os.cpu_count(),
                          "runner": res, "collate_ms_per_batch": {"plain_torch_12_tensors": t_plain * 1e3, "numpy_into_one_block": t_block * 1e3, **t_nat},
                          "what": "BulkRunner on one GPU, utterances in memory: worker-thread collate into pinned blocks, one H2D copy per batch, forward, D2H, per-utterance crop"}
print(json.dumps(out))
