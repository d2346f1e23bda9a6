#!/bin/bash
# One prefill-kernel iteration on the GPU box: parity tests touching prefill, CTA-0 phase trace, TFLOP/s.
TAG=${1:-it}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "prefill or attention or forward" 2>&1 | tail -n 3
timeout 120 python tools/trace_prefill.py --tiles 96 --out gpurun_out/${TAG}_trace.json 2>&1 | tail -n 8
for e in ${EMUS:-0 1}; do
B200ATTN_PREFILL_EMU=$e timeout 300 python tools/bench_kernels.py --what prefill --ps-only --iters 20 --out gpurun_out/${TAG}_emu${e}.json > gpurun_out/${TAG}_emu${e}.log 2>&1
python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_emu${e}.json"))
print("EMU=$e", {k: round(v["ps_tflops"], 1) for k, v in d.items() if isinstance(v, dict) and "ps_tflops" in v})
PY
done
