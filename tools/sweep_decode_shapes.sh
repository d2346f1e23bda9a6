#!/bin/bash
# GPU box: decode cases of tools/bench_kernels.py under every CTA shape of paged_decode_kernel.
out=${1:-gpurun_out/shapes}
mkdir -p $out
for shape in w4s3 w2s3 w4s2 w2s2; do
  B200ATTN_DECODE_SHAPE=$shape python tools/bench_kernels.py --what decode --out $out/decode_$shape.json > $out/decode_$shape.log 2>&1
done
python - "$out" <<'PY'
import json, sys
out = sys.argv[1]
shapes = ["w4s3", "w2s3", "w4s2", "w2s2"]
d = {s: json.load(open(f"{out}/decode_{s}.json")) for s in shapes}
print("case".ljust(34), *(s.rjust(18) for s in shapes), "   (graph-replay us | eager us)")
for k, v in d["w4s3"].items():
    if isinstance(v, dict) and v.get("kind") == "decode":
        print(k.ljust(34), *(("%7.1f | %7.1f" % (1e3 * d[s][k]["graph_ms_median"], 1e3 * d[s][k]["ours_ms_median"])).rjust(18) for s in shapes),
              "  fa2 %.1f splits %d" % (1e3 * v.get("fa2_ms_median", 0), v["num_splits"]))
PY
