#!/bin/bash
# Sweep of the stage-0 tile size (rows per CTA) x upper-stage launch shape; then a source-level capture of the upper kernel.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for cfg in "128 1 8192" "96 1 8192" "160 1 8192" "192 1 8192" "256 1 8192" "192 2 2560" "160 2 2560" "128 2 2560"; do
  set -- $cfg
  RAHT_UP_GROUPS=$2 RAHT_UP_ELEMS=$3 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --tile-rows $1 > gpurun_out/bench_sweep.json 2> gpurun_out/bench_sweep.err
  python - "$cfg" <<'PY'
import json, sys
try:
    d = json.loads(open('gpurun_out/bench_sweep.json').read().strip().splitlines()[-1])
    print(sys.argv[1], "ms_per_step", d["ms_per_step"], {k: v["ms"] for k, v in d["stages"].items() if "raht" in k})
except Exception as e:
    print(sys.argv[1], "no bench line:", e)
PY
done
bash scripts/ncu_kernel.sh "transform_upper" 2
