#!/bin/bash
# Run the prefill parity subset + TFLOP/s + CTA-0 phase trace for each variant library named in $VARIANTS
# (built by tools/build_variant.sh); "default" = the in-tree libb200attn.so.  Env per variant: NAME[:WGS[:EMU]].
TAG=${1:-var}
mkdir -p gpurun_out
for spec in ${VARIANTS:-default}; do
IFS=: read -r name w e <<< "$spec"
unset B200ATTN_LIB
[ "$name" != default ] && export B200ATTN_LIB=$PWD/nano-vllm-v1_b200/b200attn/variants/libb200attn_$name.so
export B200ATTN_PREFILL_WGS=${w:-1} B200ATTN_PREFILL_EMU=${e:-0}
echo "== $spec"
if [ -z "$NOTEST" ]; then timeout 900 python -m pytest tests -m gpu -q -x -k "prefill or planned or attention or forward or cfg4" --timeout 300 2>&1 | tail -n 2; fi
timeout 300 python tools/bench_kernels.py --what prefill --ps-only --iters 20 --out gpurun_out/${TAG}_${name}_w${w:-1}e${e:-0}.json > gpurun_out/${TAG}_${name}_w${w:-1}e${e:-0}.log 2>&1
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_${name}_w${w:-1}e${e:-0}.json"))
    print("$spec", {k: round(v["ps_tflops"], 1) for k, v in d.items() if isinstance(v, dict) and "ps_tflops" in v})
except Exception as ex:
    print("no result", ex)
PY
timeout 120 python tools/trace_prefill.py --tiles 96 --out gpurun_out/${TAG}_${name}_w${w:-1}e${e:-0}_trace.json 2>&1 | grep -v wrote | tail -n 8
done
