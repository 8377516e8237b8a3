#!/bin/bash
# Sweep of the upper-stage launch shape (channel groups x staging elements) for the stages with many tiles.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x --timeout 600 -k "raht_pipeline or scene_roundtrip or fp32_mode" > gpurun_out/test_quick.log 2>&1
echo "tests exit $?"; tail -n 2 gpurun_out/test_quick.log
for cfg in "1 8192" "1 4608" "2 4096" "2 2560" "4 2048" "4 1280" "8 1032"; do
  set -- $cfg
  RAHT_UP_GROUPS=$1 RAHT_UP_ELEMS=$2 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sweep.json 2> gpurun_out/bench_sweep.err
  python - "$cfg" <<'PY'
import json, sys
try:
    d = json.loads(open('gpurun_out/bench_sweep.json').read().strip().splitlines()[-1])
    print(sys.argv[1], "ms_per_step", d["ms_per_step"], {k: v["ms"] for k, v in d["stages"].items() if k in ("raht", "iraht")})
except Exception as e:
    print(sys.argv[1], "no bench line:", e)
PY
done
RAHT_UP_GROUPS=4 RAHT_UP_ELEMS=2048 timeout 600 python -m pytest tests -m gpu -q -x --timeout 600 -k "raht_pipeline or scene_roundtrip or fp32_mode" > gpurun_out/test_quick2.log 2>&1
echo "tests (4 x 2048) exit $?"; tail -n 2 gpurun_out/test_quick2.log
