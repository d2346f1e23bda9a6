#!/bin/bash
# Drop-in arm end to end with (1) and without (0) the peer-memory all-reduce at tensor-parallel degree $1.
# Usage: bash tools/gpu_ar_e2e.sh TP [num_seqs] [tag]
TP=${1:-2}; NS=${2:-256}; TAG=${3:-ar_e2e_tp${TP}_bs${NS}}
mkdir -p gpurun_out
for ar in 1 0; do
  B200ATTN_ALLREDUCE=$ar timeout 600 python tools/e2e_generate.py --arm dropin --size 8b --tp $TP --num-seqs $NS --in-len 128 --out-len 256 \
    --chunked-prefill --out gpurun_out/${TAG}_ar${ar}.json > gpurun_out/${TAG}_ar${ar}.log 2>&1
  echo "ar=$ar exit $?"; python -c "import json; d=json.load(open('gpurun_out/${TAG}_ar${ar}.json')); print('B200ATTN_ALLREDUCE=$ar tok/s', round(d['tok_per_s']), d['seconds_all'])"
done
