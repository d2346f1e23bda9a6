#!/bin/bash
# A/B of the persistent prefill kernel's knobs on the GPU box: parity tests touching prefill, then TFLOP/s,
# for every WGS (softmax warpgroups per query tile) x EMU (polynomial exp2 share) combination in $COMBOS.
TAG=${1:-ab}
mkdir -p gpurun_out
for c in ${COMBOS:-"1:0 2:0 2:1"}; do
w=${c%%:*}; e=${c##*:}
export B200ATTN_PREFILL_WGS=$w B200ATTN_PREFILL_EMU=$e
echo "== WGS=$w EMU=$e"
timeout 900 python -m pytest tests -m gpu -q -x -k "prefill or planned or attention or forward or cfg4" --timeout 300 2>&1 | tail -n 3
timeout 300 python tools/bench_kernels.py --what prefill --ps-only --iters 20 --out gpurun_out/${TAG}_w${w}e${e}.json > gpurun_out/${TAG}_w${w}e${e}.log 2>&1
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_w${w}e${e}.json"))
    print("WGS=$w EMU=$e", {k: round(v["ps_tflops"], 1) for k, v in d.items() if isinstance(v, dict) and "ps_tflops" in v})
except Exception as ex:
    print("no result", ex)
PY
if [ -n "$TRACE" ]; then timeout 120 python tools/trace_prefill.py --tiles 96 --out gpurun_out/${TAG}_w${w}e${e}_trace.json 2>&1 | tail -n 8; fi
done
