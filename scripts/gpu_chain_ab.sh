#!/bin/bash
# A/B of the transform launch structure (RAHT_CHAIN = 0 per-stage launches, 1 upper stages chained, 2 everything chained)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/summary_chain.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout 600 -k "golden_raht or oracle_raht or fused_quant" > gpurun_out/test_chain.log 2>&1
echo "tests (mode 1) exit $?" | tee -a gpurun_out/summary_chain.txt; tail -n 3 gpurun_out/test_chain.log | tee -a gpurun_out/summary_chain.txt
for m in 1 0 2; do
  for wl in c2 c3; do
    RAHT_CHAIN=$m timeout 600 python bench.py --steps 5 --warmup 3 --no-subrecords --no-cpu-baseline --workload $wl > gpurun_out/bench_chain_${m}_${wl}.json 2> gpurun_out/bench_chain.err
    python - "$m" "$wl" <<'PY' | tee -a gpurun_out/summary_chain.txt
import json, sys
m, wl = sys.argv[1], sys.argv[2]
try:
    d = json.loads(open(f"gpurun_out/bench_chain_{m}_{wl}.json").read().strip().splitlines()[-1])
    s = d["stages"]
    print(f"chain={m} {wl}: step {d['ms_per_step']:.3f} ms launches {d['gpu_launches']} raht {s['raht']['ms']:.4f} ({s['raht']['frac_of_peak']}) iraht {s['iraht']['ms']:.4f} ({s['iraht']['frac_of_peak']}) "
          f"raht_param {s['raht_param']['ms']:.4f} dec {s['raht_param_dec']['ms']:.4f} rlgr_enc {s['rlgr_encode']['ms']:.4f} rlgr_dec {s['rlgr_decode']['ms']:.4f}")
except Exception as e:
    print("parse failed", m, wl, e)
PY
  done
done
