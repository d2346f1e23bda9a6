#!/bin/bash
# ncu evidence for round 2 (one GPU): launch list of bench.py, --set full of the decode kernel per shape and of every byte mover.
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 1 --no-tok-per-s --no-gpu-baseline --no-cpu-baseline > gpurun_out/r2_bench_short.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_bench_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-tok-per-s --no-gpu-baseline --no-cpu-baseline > gpurun_out/r2_bench_ncu.log 2>&1
echo "launch list rc=$?"
timeout 600 python tools/prof_decode.py > gpurun_out/r2_prof_decode.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:paged_decode -o gpurun_out/r2_decode python tools/prof_decode.py > gpurun_out/r2_decode_ncu.log 2>&1
echo "decode rc=$?"
BK_WARM=1 timeout 600 python tools/bench_kernels.py --what pointwise --iters 1 --out gpurun_out/r2_pw_tmp.json > /dev/null 2>&1 && \
BK_WARM=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'kv_scatter|qk_norm_rope|rmsnorm|silu_and_mul' -o gpurun_out/r2_pointwise \
    python tools/bench_kernels.py --what pointwise --iters 1 --out gpurun_out/r2_pw_tmp.json > gpurun_out/r2_pointwise_ncu.log 2>&1
echo "pointwise rc=$?"
ls -la gpurun_out | tail -12
