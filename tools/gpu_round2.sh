#!/bin/bash
# One GPU-box visit (round 2): smoke, GPU tests, bench (both arms), kernel table.
# Usage (from the repo root on the box): bash tools/gpu_round2.sh [tag]
TAG=${1:-r2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/${TAG}_gpu.csv 2>&1
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/${TAG}_smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/${TAG}_smoke.log
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 --durations=15 > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 50 --warmup 3 > gpurun_out/${TAG}_bench.log 2>&1
echo "bench exit $?" >> gpurun_out/${TAG}_bench.log
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.log 2>&1
echo "bench ref exit $?" >> gpurun_out/${TAG}_bench_ref.log
timeout 600 python tools/bench_kernels.py --out gpurun_out/${TAG}_kernels.json > gpurun_out/${TAG}_kernels.log 2>&1
echo "kernels exit $?" >> gpurun_out/${TAG}_kernels.log
tail -n 4 gpurun_out/${TAG}_smoke.log gpurun_out/${TAG}_pytest.log gpurun_out/${TAG}_bench.log gpurun_out/${TAG}_bench_ref.log gpurun_out/${TAG}_kernels.log
