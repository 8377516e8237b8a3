python -m pytest tests -m gpu -q -x -k "rlgr or container or sync_index or scene_roundtrip" 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dec.json 2> gpurun_out/bench_dec.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_dec.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["e2e"]["value"], {k:v["ms"] for k,v in d["stages"].items()}, d["container"])
PY
tail -3 gpurun_out/bench_dec.err
