"""Where the host time of bulk generation goes: cProfile of BulkRunner.run (main thread) on a slice of the cfg3 stream.
    python tools/bulk_profile.py [utterances]"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from daft_exprt_b200 import bulk, synthetic  # noqa: E402
from daft_exprt_b200.hparams import default_hparams  # noqa: E402
from daft_exprt_b200.model import DaftExprt  # noqa: E402
from daft_exprt_b200.weights import make_state_dict  # noqa: E402

n_utt = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
hp = default_hparams()
model = DaftExprt(hp, mode="bf16").cuda()
model.load_state_dict(make_state_dict(hp, 1234))
in_lens, out_lens = synthetic.stream_lengths(hp, n_utt, seed=43)
batches, mine = bulk.plan_shards(in_lens, out_lens, 64, 1)
utts = [synthetic.utterance(hp, g, seed=43) for g in range(n_utt)]
stream = {f: [u[f] for u in utts] for f in synthetic.UtteranceStream.FIELDS}
table = bulk.PointerTable(stream)
kw = dict(workers=2, slots=6, collate_threads=max(1, min(4, len(os.sched_getaffinity(0)) // 2)))
ids = list(range(len(batches)))
for _ in range(2):
    r = bulk.BulkRunner(model, sink=lambda i, mel: None, **kw)
    t0 = time.perf_counter()
    frames = r.run(table, batches, ids)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
print(f"{len(batches)} batches, {frames} frames, {dt * 1e3 / len(batches):.3f} ms per batch, {frames / dt / 1e6:.2f} M frames/s")
# device-only time of the same batches (inputs resident, no copies): the floor
r = bulk.BulkRunner(model, sink=lambda i, mel: None, **kw)
pr = cProfile.Profile()
pr.enable()
r.run(table, batches, ids)
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(18)
