timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_t.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["e2e"]["value"], {k:v["ms"] for k,v in d["stages"].items()})
PY
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload c3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_c3.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["e2e"]["value"], {k:v["ms"] for k,v in d["stages"].items()})
PY
