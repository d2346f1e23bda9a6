#!/bin/bash
# (CTA shape x split count) sweep of the decode kernel on the sharded shapes; one process per point.
for shape in w4s3 w2s3 w4s2 w2s2; do for s in ${SPLITS:-auto 1 2 3 4 6 8}; do
  if [ $s = auto ]; then B200ATTN_DECODE_SHAPE=$shape python tools/sweep_decode_tp.py "$@"; else B200ATTN_DECODE_SHAPE=$shape B200ATTN_DECODE_SPLITS=$s python tools/sweep_decode_tp.py "$@"; fi
done; done 2>&1 | grep -v Warning | tee gpurun_out/r2_sweep_decode_tp.txt
