#!/bin/bash
# One GPU-box visit: smoke, GPU tests, bench, then the ncu launch list of the same bench command.
# Usage (from the repo root on the box): bash tools/gpu_round.sh [tag]
TAG=${1:-r1}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/${TAG}_gpu.csv 2>&1
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/${TAG}_smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/${TAG}_smoke.log
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.log 2>&1
echo "bench exit $?" >> gpurun_out/${TAG}_bench.log
tail -n 3 gpurun_out/${TAG}_smoke.log gpurun_out/${TAG}_pytest.log gpurun_out/${TAG}_bench.log
