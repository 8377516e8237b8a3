#!/bin/bash
# The end-of-round evidence run without the `--set full` capture: whole GPU suite, smoke, driver's bench command, launch list.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/summary.txt
say() { echo "$@" | tee -a gpurun_out/summary.txt; }
say "=== all gpu tests in one process"
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/test_all.log 2>&1; say "exit $?"; tail -n 3 gpurun_out/test_all.log | tee -a gpurun_out/summary.txt
say "=== smoke"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; say "exit $?"; tail -n 2 gpurun_out/smoke.log | tee -a gpurun_out/summary.txt
say "=== bench (driver's command)"
timeout 1200 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; say "exit $?"
tail -c 1200 gpurun_out/bench.json | tee -a gpurun_out/summary.txt; tail -n 3 gpurun_out/bench.err | tee -a gpurun_out/summary.txt
say "=== bench again, headline only"
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-subrecords > gpurun_out/bench_nosub.json 2> gpurun_out/bench_nosub.err; say "exit $?"
say "=== ncu launch list"
timeout 300 python scripts/step_once.py > gpurun_out/step_once.log 2>&1 && bash scripts/ncu_launches.sh | tee -a gpurun_out/summary.txt
python scripts/ncu_summary.py gpurun_out/launches.csv > gpurun_out/launches.txt 2>&1; head -n 16 gpurun_out/launches.txt
python - <<'PY' | tee -a gpurun_out/summary.txt
import json
for f in ("gpurun_out/bench.json", "gpurun_out/bench_nosub.json"):
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, "ms_per_step", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "launches", d["gpu_launches"], "parity", d["parity"]["streams_equal"], d["parity"]["decoded_attributes_equal"])
    print({k: round(v["ms"], 4) for k, v in d["stages"].items()})
PY
