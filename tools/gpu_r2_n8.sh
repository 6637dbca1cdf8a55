#!/bin/bash
# end of round 2: the default bench on 8 GPUs of one box, launched as the driver launches it
mkdir -p gpurun_out
timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2g_bench_8gpu.json 2> gpurun_out/r2g_bench_8gpu.err; echo "rc=$?"
tail -c 300 gpurun_out/r2g_bench_8gpu.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2g_bench_8gpu.json").read().strip().splitlines()[-1])
c=d.get("configs") or {}
print(d["n_gpus"], "value", round(d["value"]/1e6,2), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]/1e6,2), "cfg3", c.get("cfg3",{}).get("frames_s"), c.get("cfg3",{}).get("per_rank_seconds"), "cfg5", {k: c.get("cfg5",{}).get(k) for k in ("step_ms","all_reduce_ms","frames_s")}, "eval", c.get("cfg5",{}).get("eval_mode",{}).get("step_ms"))
PY
