#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {  # label, env...
  label="$1"; shift
  env "$@" timeout 600 python bench.py --steps 20 --warmup 5 --no-subrecords > "gpurun_out/d_$label.json" 2> "gpurun_out/d_$label.err"; echo "bench $label ($*) exit $?"
  python - "$label" <<'PY'
import json, sys
d = json.loads(open(f'gpurun_out/d_{sys.argv[1]}.json').read().strip().splitlines()[-1])
p = d.get("parity") or {}
print("ms_per_step", d["ms_per_step"], "e2e", round(d["e2e"]["value"]/1e6,1), "parity", p.get("streams_equal"), p.get("keys_equal"), p.get("decoded_attributes_equal"))
print({k: round(v["ms"], 4) for k, v in d["stages"].items()})
PY
}
run prio11 RAHT_DEC_SIDE=1 RAHT_ENC_SIDE=1
run prio11_m3 RAHT_DEC_SIDE=1 RAHT_ENC_SIDE=1 RAHT_MERGE_CTAS=3
run noprio11 RAHT_DEC_SIDE=1 RAHT_ENC_SIDE=1 RAHT_SIDE_PRIO=0
run m3_only RAHT_MERGE_CTAS=3
