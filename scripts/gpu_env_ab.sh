#!/bin/bash
# A/B of an environment knob of the library: bash scripts/gpu_env_ab.sh VAR val1 val2 ...  (tests once with the default, then a bench per value)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
VAR="$1"; shift
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "${AB_TESTS:-rlgr or scene or container or sharded}" > gpurun_out/test_ab.log 2>&1
echo "tests exit $?: $(tail -n 1 gpurun_out/test_ab.log)"
for v in "$@"; do
  env "$VAR=$v" timeout 600 python bench.py --steps 20 --warmup 5 --no-subrecords > "gpurun_out/ab_$v.json" 2> "gpurun_out/ab_$v.err"; echo "bench $VAR=$v exit $?"
  python - "$v" <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/ab_{sys.argv[1]}.json').read().strip().splitlines()[-1])
p = d.get("parity") or {}
print("ms_per_step", d["ms_per_step"], "launches", d["gpu_launches"], "parity", p.get("streams_equal"), p.get("keys_equal"), p.get("decoded_attributes_equal"), "sparse", (d.get("sparse") or {}).get("ms_per_step"))
print({k: round(v["ms"], 4) for k, v in d["stages"].items()})
PY
done
