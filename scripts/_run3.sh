timeout 600 python -m pytest tests -m gpu -q -x -k "rlgr or container or sync_index or scene_roundtrip or sharded" 2>&1 | tail -3
RLGR_PROBE_SPLIT=1 timeout 300 python scripts/rlgr_probe.py 2>&1 | tail -8
timeout 300 python scripts/rlgr_probe.py 6000000 0.015625 2>&1 | tail -2
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_h.json 2> gpurun_out/bench_h.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_h.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["e2e"]["value"], {k:v["ms"] for k,v in d["stages"].items()})
PY
