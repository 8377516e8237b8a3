#!/bin/bash
# ncu launch list (gpu__time_duration per launch) of two encode + decode steps -> gpurun_out/launches.csv
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python scripts/step_once.py "$@" > gpurun_out/ncu.log 2>&1
echo "ncu exit $?"
