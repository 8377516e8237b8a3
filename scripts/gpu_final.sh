#!/bin/bash
# End-of-round evidence run on one B200: the whole GPU suite in one process (as the driver runs it), smoke(), the default
# bench line, the reference arm (2 steps), then -- each only after its plain run exited 0 -- the ncu launch list and the
# `--set full` capture of two steps.  Logs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
: > gpurun_out/summary.txt
say() { echo "$@" | tee -a gpurun_out/summary.txt; }
say "=== all gpu tests in one process"
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/test_all.log 2>&1; say "exit $?"; tail -n 3 gpurun_out/test_all.log | tee -a gpurun_out/summary.txt
say "=== smoke"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; say "exit $?"; tail -n 2 gpurun_out/smoke.log | tee -a gpurun_out/summary.txt
say "=== bench (driver's command)"
timeout 1200 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; say "exit $?"
tail -c 1500 gpurun_out/bench.json | tee -a gpurun_out/summary.txt; tail -n 3 gpurun_out/bench.err | tee -a gpurun_out/summary.txt
say "=== reference arm (2 steps)"
timeout 900 python bench.py --impl reference --gpus 1 --steps 2 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; say "exit $?"
tail -c 600 gpurun_out/bench_ref.json | tee -a gpurun_out/summary.txt
say "=== ncu launch list"
timeout 300 python scripts/step_once.py > gpurun_out/step_once.log 2>&1 && bash scripts/ncu_launches.sh | tee -a gpurun_out/summary.txt
python scripts/ncu_summary.py gpurun_out/launches.csv > gpurun_out/launches.txt 2>&1; head -n 30 gpurun_out/launches.txt
say "=== ncu --set full"
bash scripts/ncu_full.sh | tail -n 3
python scripts/ncu_full_summary.py gpurun_out/full_raw.csv --json gpurun_out/ncu_full.json > gpurun_out/ncu_full.txt 2>&1; head -n 12 gpurun_out/ncu_full.txt
