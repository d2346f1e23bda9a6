#!/bin/bash
# A/B sweep of the persistent prefill kernel over B200ATTN_PREFILL_EMU (share of exponentials taken by
# the FMA-pipe polynomial instead of MUFU.ex2).  GPU box only.  Usage: bash tools/sweep_prefill_emu.sh [tag]
TAG=${1:-emu}
mkdir -p gpurun_out
for e in 0 1 2 3 4; do
  B200ATTN_PREFILL_EMU=$e timeout 300 python tools/bench_kernels.py --what prefill --ps-only --iters 20 \
    --out gpurun_out/${TAG}_emu${e}.json > gpurun_out/${TAG}_emu${e}.log 2>&1
  echo "EMU=$e exit $?"
  python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_emu${e}.json"))
print({k: round(v["ps_tflops"], 1) for k, v in d.items() if isinstance(v, dict) and "ps_tflops" in v})
PY
done
