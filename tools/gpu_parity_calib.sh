#!/bin/bash
# Calibration of the engine-level greedy parity test (tests/test_gpu_engine.py): reference arm, drop-in arm free-running,
# drop-in arm teacher-forced, for a given number of layers.  Writes gpurun_out/calib_L<layers>_{free,forced}.json.
set -u
L=${1:-4}
OUT=gpurun_out
COMMON="--size 0.6b --layers $L --num-seqs 32 --in-len 128 --out-len 64 --chunked-prefill --max-model-len 1024 --max-num-batched-tokens 1000 --gpu-memory-utilization 0.3 --greedy --fp32-logits"
python tools/e2e_generate.py --arm reference $COMMON --out /tmp/calib_ref_$L.json > $OUT/calib_L${L}_ref.log 2>&1
python tools/e2e_generate.py --arm dropin $COMMON --out /tmp/calib_our_$L.json > $OUT/calib_L${L}_our.log 2>&1
python tools/e2e_generate.py --arm dropin $COMMON --force-tokens /tmp/calib_ref_$L.json --out /tmp/calib_tf_$L.json > $OUT/calib_L${L}_tf.log 2>&1
python tools/e2e_generate.py --compare-free /tmp/calib_ref_$L.json /tmp/calib_our_$L.json --tie-margin ${2:-0.03} --out $OUT/calib_L${L}_free.json
python tools/e2e_generate.py --compare /tmp/calib_ref_$L.json /tmp/calib_tf_$L.json --tie-margin ${2:-0.03} --out $OUT/calib_L${L}_forced.json
