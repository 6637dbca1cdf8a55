#!/bin/bash
# timing experiments on the fused FFN kernels: knob 7 (0 = CTA pair, 1 = single CTA) x knob 8 bits
# (1 = no weight copies, 2 = no MMAs, 4 = no Hs stores).  Results are wrong for 8 != 0; only the kernel time matters.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for k7 in 0 1; do for k8 in 0 1 2 3 4 5; do
  DX_DEBUG="7=$k7,8=$k8" timeout 120 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('knob7=$k7 knob8=$k8 ffn_ms=%.4f step_ms=%.4f' % (d['kernels_ms_per_step']['tc_ffn_fused'], d['ms_per_step']))"
done; done | tee gpurun_out/ffn_knobs.txt
