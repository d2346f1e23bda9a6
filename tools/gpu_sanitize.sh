#!/bin/bash
# One compute-sanitizer tool over the small-shape kernel parity tests (every kernel of csrc/ through the C-ABI):
#   gpurun --timeout 1500 -- 'bash tools/gpu_sanitize.sh memcheck'      (or racecheck | initcheck | synccheck)
# One tool per GPU call; the log and the pytest tail land in gpurun_out/r2_sanitizer_<tool>.{log,txt}.
set -u
TOOL=${1:-memcheck}
OUT=gpurun_out
mkdir -p $OUT
SEL='(scatter or decode_ragged or decode_minus or decode_cuda_graph or prefill_unpaged or prefill_paged or planned_paged or planned_unpaged or planned_chunked or pointwise or fused_norm_rope or decode_many or peer_memory or single_token or tickets or nvls or cfg4 or prefix_sharing or eager_decode) and not full_size and not golden'
EXTRA=""
[ "$TOOL" = initcheck ] && EXTRA="--track-unused-memory no"
[ "$TOOL" = memcheck ] && EXTRA="--leak-check no"
timeout ${SAN_TIMEOUT:-1300} compute-sanitizer --tool $TOOL $EXTRA --error-exitcode 9 --print-limit 40 \
    --log-file $OUT/r2_sanitizer_$TOOL.log \
    python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "$SEL" -p no:cacheprovider > $OUT/r2_sanitizer_$TOOL.txt 2>&1
echo "rc=$?" >> $OUT/r2_sanitizer_$TOOL.txt
tail -5 $OUT/r2_sanitizer_$TOOL.txt
grep -c "=========" $OUT/r2_sanitizer_$TOOL.log
tail -8 $OUT/r2_sanitizer_$TOOL.log
