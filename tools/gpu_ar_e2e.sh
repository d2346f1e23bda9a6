#!/bin/bash
# TP2 end to end with (1) and without (0) the peer-memory all-reduce, then the parity/latency check.  2-GPU box.
for ar in 1 0; do
  B200ATTN_ALLREDUCE=$ar timeout 600 python tools/e2e_generate.py --arm dropin --size 8b --tp 2 --num-seqs 4 --in-len 128 --out-len 256 --chunked-prefill --out gpurun_out/ar_e2e_tp2_ar${ar}.json > gpurun_out/ar_e2e_tp2_ar${ar}.log 2>&1
  echo "ar=$ar exit $?"; tail -n 1 gpurun_out/ar_e2e_tp2_ar${ar}.log | cut -c1-400
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/ar_check.py 2>&1 | grep -v Warn | tail -7
