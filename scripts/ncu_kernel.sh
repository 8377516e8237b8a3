#!/bin/bash
# `ncu --set full` (+ source) of the kernels matching a regex during two encode + decode steps.
# usage: bash scripts/ncu_kernel.sh <kernel-regex> [count] [step_once args...] -> gpurun_out/k.ncu-rep, k_raw.csv, k_src.csv
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
RE="$1"; CNT="${2:-2}"; shift; shift
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$RE" -c "$CNT" -f -o gpurun_out/k python scripts/step_once.py "$@" > gpurun_out/ncu_k.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/k.ncu-rep --page raw --csv > gpurun_out/k_raw.csv 2>/dev/null
ncu -i gpurun_out/k.ncu-rep --page source --csv > gpurun_out/k_src.csv 2>/dev/null
if [ "$(stat -c %s gpurun_out/k.ncu-rep 2>/dev/null || echo 0)" -gt 30000000 ]; then rm -f gpurun_out/k.ncu-rep; fi
