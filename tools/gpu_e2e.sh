#!/bin/bash
# cfg2 end to end on one B200: reference arm, then drop-in arm, greedy parity + throughput.
# Usage: bash tools/gpu_e2e.sh [tag] [extra e2e_generate args]
TAG=${1:-e2e}; shift
mkdir -p gpurun_out
COMMON="--size 0.6b --num-seqs 256 --in-len 128 --out-len 256 --chunked-prefill $@"
timeout 900 python tools/e2e_generate.py --arm reference $COMMON --greedy --out gpurun_out/${TAG}_reference_greedy.json > gpurun_out/${TAG}_reference_greedy.log 2>&1
echo "reference greedy exit $?"; tail -n 1 gpurun_out/${TAG}_reference_greedy.log | cut -c1-400
# drop-in arm is teacher-forced with the reference's tokens: every decision is compared on an identical context
timeout 900 python tools/e2e_generate.py --arm dropin $COMMON --greedy --force-tokens gpurun_out/${TAG}_reference_greedy.json --out gpurun_out/${TAG}_dropin_greedy.json > gpurun_out/${TAG}_dropin_greedy.log 2>&1
echo "dropin greedy (teacher-forced) exit $?"; tail -n 1 gpurun_out/${TAG}_dropin_greedy.log | cut -c1-400
python tools/e2e_generate.py --compare gpurun_out/${TAG}_reference_greedy.json gpurun_out/${TAG}_dropin_greedy.json --out gpurun_out/${TAG}_parity.json
for arm in reference dropin; do
  timeout 900 python tools/e2e_generate.py --arm $arm $COMMON --out gpurun_out/${TAG}_${arm}.json > gpurun_out/${TAG}_${arm}.log 2>&1
  echo "$arm sampled exit $?"; tail -n 1 gpurun_out/${TAG}_${arm}.log | cut -c1-600
done
