#!/bin/bash
# Quick GPU loop: the test groups named in $1 (pytest -k expression, default: rlgr or scene), then a short bench.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
K="${1:-rlgr or scene}"
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
