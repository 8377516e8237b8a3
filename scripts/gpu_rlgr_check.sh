#!/bin/bash
# RLGR encoder iteration: parity of the coder on the GPU, then the probe (encode alone) for a few warm-up lengths.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/summary_rlgr.txt
echo "=== tests" | tee -a gpurun_out/summary_rlgr.txt
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_container.py -m gpu -q -x --timeout 900 -k "rlgr or golden_raht or oracle_raht or container or scene_roundtrip" > gpurun_out/test_rlgr.log 2>&1
echo "exit $?" | tee -a gpurun_out/summary_rlgr.txt; tail -n 12 gpurun_out/test_rlgr.log | tee -a gpurun_out/summary_rlgr.txt
timeout 1200 python -m pytest tests/test_gpu_scale_parity.py -m gpu -q -x --timeout 900 -k "c2 or without_index" > gpurun_out/test_scale_c2.log 2>&1
echo "exit $?" | tee -a gpurun_out/summary_rlgr.txt; tail -n 12 gpurun_out/test_scale_c2.log | tee -a gpurun_out/summary_rlgr.txt
for w in ${WARMS:-48 16 32 64 128}; do
  echo "=== probe warm=$w" | tee -a gpurun_out/summary_rlgr.txt
  RAHT_RLGR_WARM=$w timeout 300 python scripts/rlgr_probe.py 1000000 0.015625 2>&1 | tail -n 2 | tee -a gpurun_out/summary_rlgr.txt
done
echo "=== probe sparse step 0.25" | tee -a gpurun_out/summary_rlgr.txt
timeout 300 python scripts/rlgr_probe.py 1000000 0.25 2>&1 | tail -n 2 | tee -a gpurun_out/summary_rlgr.txt
echo "=== probe 6M" | tee -a gpurun_out/summary_rlgr.txt
RLGR_PROBE_J=16 timeout 600 python scripts/rlgr_probe.py 6000000 0.015625 2>&1 | tail -n 2 | tee -a gpurun_out/summary_rlgr.txt
echo "=== bench" | tee -a gpurun_out/summary_rlgr.txt
timeout 900 python bench.py --steps 5 --warmup 3 --no-subrecords > gpurun_out/bench_rlgr.json 2> gpurun_out/bench_rlgr.err; echo "exit $?" | tee -a gpurun_out/summary_rlgr.txt
python - <<'PY' | tee -a gpurun_out/summary_rlgr.txt
import json
try:
    d = json.loads(open("gpurun_out/bench_rlgr.json").read().strip().splitlines()[-1])
    print("ms_per_step", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "launches", d["gpu_launches"], "parity", d["parity"])
    for k, v in d["stages"].items():
        print(f"  {k:20s} {v['ms']:.4f} ms  {v['frac_of_peak']}")
except Exception as e:
    print("bench parse failed", e)
PY
tail -n 5 gpurun_out/bench_rlgr.err | tee -a gpurun_out/summary_rlgr.txt
