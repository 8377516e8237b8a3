#!/bin/bash
# quick iteration: a test selection, then the bench headline (no sub-records)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/summary_quick.txt
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -k "${1:-scene}" > gpurun_out/test_quick.log 2>&1
echo "tests exit $?" | tee -a gpurun_out/summary_quick.txt; tail -n 15 gpurun_out/test_quick.log | tee -a gpurun_out/summary_quick.txt
timeout 900 python bench.py --steps 5 --warmup 3 --no-subrecords ${BENCH_ARGS} > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit $?" | tee -a gpurun_out/summary_quick.txt
python - <<'PY' | tee -a gpurun_out/summary_quick.txt
import json
try:
    d = json.loads(open("gpurun_out/bench_quick.json").read().strip().splitlines()[-1])
    print("ms_per_step", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "attr", d["e2e"]["with_attributes"]["value"], "launches", d["gpu_launches"], "mallocs", d["cuda_mallocs_in_timed_region"])
    print("parity", d["parity"])
    for k, v in d["stages"].items():
        print(f"  {k:20s} {v['ms']:.4f} ms  {v['frac_of_peak']}")
except Exception as e:
    print("bench parse failed", e)
PY
tail -n 8 gpurun_out/bench_quick.err | tee -a gpurun_out/summary_quick.txt
