#!/bin/bash
# Qwen3-8B tokens/s through the untouched LLM.generate at tensor-parallel degree $1, reference arm and
# drop-in arm back to back on the same box.  Usage: bash tools/gpu_e2e_tp.sh TP [tag]
TP=${1:-2}; TAG=${2:-e2e_tp${TP}}
mkdir -p gpurun_out
for wl in "--in-len 128 --out-len 256 --chunked-prefill" "--in-len 1024 --out-len 256"; do
  name=$(echo $wl | tr -d ' -' )
  for arm in reference dropin; do
    timeout 900 python tools/e2e_generate.py --arm $arm --size 8b --tp $TP --num-seqs 256 $wl \
      --out gpurun_out/${TAG}_${name}_${arm}.json > gpurun_out/${TAG}_${name}_${arm}.log 2>&1
    echo "tp$TP $name $arm exit $?"; tail -n 1 gpurun_out/${TAG}_${name}_${arm}.log | cut -c1-300
  done
done
python - <<PY
import glob, json
out = []
for f in sorted(glob.glob("gpurun_out/${TAG}_*_*.json")):
    try:
        d = json.load(open(f)); out.append(d); print(f, d.get("arm"), d.get("tok_per_s"))
    except Exception as e:
        print(f, "unreadable", e)
json.dump(out, open("gpurun_out/${TAG}.json", "w"), indent=1)
PY
