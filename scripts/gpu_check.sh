#!/bin/bash
# Runs the GPU test groups in separate processes (a CUDA fault in one group must not poison the rest),
# then smoke() and a short bench.  Logs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() {  # name, pytest -k expression
  echo "=== $1" | tee -a gpurun_out/summary.txt
  timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "$2" > "gpurun_out/test_$1.log" 2>&1
  echo "exit $?" | tee -a gpurun_out/summary.txt
  tail -n 3 "gpurun_out/test_$1.log" | tee -a gpurun_out/summary.txt
}
: > gpurun_out/summary.txt
run sort "pinning or radix_sort"
run voxelize "golden_voxelize or merge_errors"
run voxmerge "oracle_voxelize_merge"
run golden_raht "golden_raht_pipeline"
run oracle_raht "oracle_raht_pipeline"
run rlgr "rlgr"
run misc "plain_lists or raht_param_errors or fp32_mode"
run scene "scene_roundtrip or scene_pipeline or fused or per_attribute or scene_with"
run sharded "sharded"
run container "container or geometry_symbols or sync_index_matches"
run reference_live "reference_kernel or reference_python or reference_voxelizer"
run scale_parity "bench_scene_is_bit_exact or without_index_decode_at_scale"
echo "=== smoke" | tee -a gpurun_out/summary.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "exit $?" | tee -a gpurun_out/summary.txt
tail -n 2 gpurun_out/smoke.log | tee -a gpurun_out/summary.txt
echo "=== bench" | tee -a gpurun_out/summary.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "exit $?" | tee -a gpurun_out/summary.txt
tail -c 3000 gpurun_out/bench.json | tee -a gpurun_out/summary.txt
tail -n 5 gpurun_out/bench.err | tee -a gpurun_out/summary.txt
echo "=== all gpu tests in one process (as the driver runs them)" | tee -a gpurun_out/summary.txt
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/test_all.log 2>&1; echo "exit $?" | tee -a gpurun_out/summary.txt
tail -n 3 gpurun_out/test_all.log | tee -a gpurun_out/summary.txt
