#!/bin/bash
# side-stream knobs of the scene-level calls: bench per setting
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 -k "rlgr or scene_roundtrip or scene_pipeline" > gpurun_out/test_ab.log 2>&1
echo "tests exit $?: $(tail -n 1 gpurun_out/test_ab.log)"
for cfg in "0 0" "1 0" "0 1" "1 1"; do
  set -- $cfg
  RAHT_DEC_SIDE=$1 RAHT_ENC_SIDE=$2 timeout 600 python bench.py --steps 20 --warmup 5 --no-subrecords > "gpurun_out/side_$1$2.json" 2> "gpurun_out/side_$1$2.err"; echo "bench dec_side=$1 enc_side=$2 exit $?"
  python - "$1$2" <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/side_{sys.argv[1]}.json').read().strip().splitlines()[-1])
p = d.get("parity") or {}
print("ms_per_step", d["ms_per_step"], "e2e", round(d["e2e"]["value"]/1e6,1), "parity", p.get("streams_equal"), p.get("keys_equal"), p.get("decoded_attributes_equal"))
print({k: round(v["ms"], 4) for k, v in d["stages"].items()})
PY
done
