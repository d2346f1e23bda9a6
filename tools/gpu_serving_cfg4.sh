#!/bin/bash
# BASELINE.json configs[3] end to end: Qwen3-14B (random-init), TP=2, request rate 10, 1024 requests, chunked prefill,
# max-num-batched-tokens 1024 and 8192, both arms.   gpurun --gpus 2 --timeout 1500 -- 'bash tools/gpu_serving_cfg4.sh'
mkdir -p gpurun_out
for mnbt in ${MNBT:-1024 8192}; do
for arm in reference dropin; do
  timeout ${RUN_TIMEOUT:-420} python tools/serving_bench.py --arm $arm --size 14b --tp 2 --request-rate 10 --num-requests ${NREQ:-1024} \
      --max-num-batched-tokens $mnbt --max-num-seqs 1024 --chunked-prefill --out gpurun_out/r2_serving_14b_tp2_mnbt${mnbt}_$arm.json \
      > gpurun_out/r2_serving_14b_tp2_mnbt${mnbt}_$arm.log 2>&1
  echo "mnbt $mnbt $arm rc=$?"; tail -n 1 gpurun_out/r2_serving_14b_tp2_mnbt${mnbt}_$arm.log | cut -c1-700
done
done
