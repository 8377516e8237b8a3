#!/bin/bash
# round-2 call B: whole GPU suite, then the bench with the RLGR staging on / off
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/test_all.log 2>&1; echo "all tests exit $?"; tail -n 5 gpurun_out/test_all.log
for s in 1 0; do
  RAHT_RLGR_STAGE=$s timeout 600 python bench.py --steps 20 --warmup 5 --no-subrecords > gpurun_out/stage_$s.json 2> gpurun_out/stage_$s.err; echo "bench stage=$s exit $?"
  python - $s <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/stage_{sys.argv[1]}.json').read().strip().splitlines()[-1])
p = d.get("parity") or {}
print("ms_per_step", d["ms_per_step"], "parity", p.get("streams_equal"), p.get("keys_equal"), p.get("decoded_attributes_equal"))
print({k: round(v["ms"], 4) for k, v in d["stages"].items()})
PY
done
bash scripts/ncu_launches.sh > /dev/null 2>&1
python scripts/ncu_summary.py gpurun_out/launches.csv > gpurun_out/launches.txt 2>&1; grep -i "rlgr" gpurun_out/launches.txt
