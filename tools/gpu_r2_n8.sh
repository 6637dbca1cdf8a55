#!/bin/bash
# round 2: the default bench on 8 GPUs of one box (one process per GPU), then the e2e part alone without host binding
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo8.txt 2>&1; lscpu | grep -E "^CPU\(s\)|NUMA|Model name|Socket" >> gpurun_out/r2_topo8.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench_8gpu.json 2> gpurun_out/r2_bench_8gpu.err; echo "rc=$?"
tail -c 400 gpurun_out/r2_bench_8gpu.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 --no-bind --no-configs --no-cpu-baseline > gpurun_out/r2_bench_8gpu_nobind.json 2> gpurun_out/r2_bench_8gpu_nobind.err; echo "rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_2gpu.json 2> gpurun_out/r2_bench_2gpu.err; echo "rc=$?"
python - <<'PY'
import json
for f in ("r2_bench_8gpu","r2_bench_8gpu_nobind","r2_bench_2gpu"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["n_gpus"], "value", round(d["value"]/1e6,2), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]/1e6,2), round(d["e2e"]["ms_per_step"],4), d["e2e"].get("host_binding",{}).get("numa_node"), (d.get("configs") or {}).get("cfg3",{}).get("frames_s"), (d.get("configs") or {}).get("cfg5",{}).get("step_ms"))
    except Exception as e:
        print(f, "ERR", e)
PY
