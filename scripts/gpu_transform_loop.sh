#!/bin/bash
# Transform-kernel loop: parity tests that touch the transform, a short bench, the ncu launch list of one step and
# (optionally, NCU_SRC=1) a `--set full` source capture of the transform kernels.  Logs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
K="${1:-raht_pipeline or scene_roundtrip or fused or fp32_mode or plain_lists}"
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "$K" > gpurun_out/test_quick.log 2>&1
echo "tests exit $?"; tail -n 4 gpurun_out/test_quick.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline ${BENCH_ARGS} > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
echo "bench exit $?"; tail -n 3 gpurun_out/bench_quick.err
python - <<'PY'
import json
try:
    d = json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
    print("ms_per_step", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "launches", d["gpu_launches"])
    print({k: v["ms"] for k, v in d["stages"].items()})
except Exception as e:
    print("no bench line:", e)
PY
bash scripts/ncu_launches.sh
python scripts/ncu_summary.py gpurun_out/launches.csv > gpurun_out/launches.txt 2>&1
grep -i "transform\|sum" gpurun_out/launches.txt
grep "transform" gpurun_out/launches.csv | awk -F'","' '{print $5, $NF}' | tail -n 14
if [ -n "$NCU_SRC" ]; then bash scripts/ncu_kernel.sh "transform_" 14; fi
