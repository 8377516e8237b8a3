#!/bin/bash
# A/B of library variants (scripts/build_variants.sh -> scratch/variants/<name>/libraht_b200.so) on the GPU box: for every
# variant the product library is swapped in the box's scratch copy, the transform / scene parity tests run, then a bench
# without sub-records (the bench line carries its own bit-exact `parity` record against the CPU checker).
# usage: bash scripts/gpu_variants.sh base name1 name2 ...   ("base" = the library as built in-tree)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
LIB=raht-3dgs-codec_b200/libraht_b200.so
cp "$LIB" /tmp/libraht_base.so
: > gpurun_out/variants.txt
for v in "$@"; do
  if [ "$v" = base ]; then cp /tmp/libraht_base.so "$LIB"; else cp "scratch/variants/$v/libraht_b200.so" "$LIB"; fi
  echo "=== $v" | tee -a gpurun_out/variants.txt
  timeout 600 python -m pytest tests -m gpu -q -x --timeout 300 -k "${VARIANT_TESTS:-golden_raht_pipeline or scene_roundtrip or merge or sharded}" > "gpurun_out/variant_$v.test.log" 2>&1
  echo "tests exit $?: $(tail -n 1 gpurun_out/variant_$v.test.log)" | tee -a gpurun_out/variants.txt
  timeout 600 python bench.py --steps 20 --warmup 5 --no-subrecords > "gpurun_out/variant_$v.json" 2> "gpurun_out/variant_$v.err"
  echo "bench exit $?" | tee -a gpurun_out/variants.txt
  python - "$v" <<'PY' | tee -a gpurun_out/variants.txt
import json, sys
try:
    d = json.loads(open(f'gpurun_out/variant_{sys.argv[1]}.json').read().strip().splitlines()[-1])
    p = d.get("parity") or {}
    print("ms_per_step", d["ms_per_step"], "e2e", round(d["e2e"]["value"] / 1e6, 1), "launches", d["gpu_launches"],
          "parity", p.get("streams_equal"), p.get("keys_equal"), p.get("decoded_attributes_equal"))
    print({k: round(v["ms"], 4) for k, v in d["stages"].items()})
except Exception as e:
    print("no bench line:", e)
PY
done
cp /tmp/libraht_base.so "$LIB"
