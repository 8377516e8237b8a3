#!/bin/bash
# round-2 call A: the whole GPU suite in one process on the current build, then the variants
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/test_all.log 2>&1; echo "all tests exit $?"; tail -n 3 gpurun_out/test_all.log
bash scripts/gpu_variants.sh "$@"
