#!/bin/bash
# One `ncu --set full` capture of every kernel of two encode + decode steps (scripts/step_once.py; run only after it
# exited 0 without ncu).  usage: bash scripts/ncu_full.sh [kernel-regex] [count]  -> gpurun_out/full_raw.csv (+ full.ncu-rep if small)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
RE="${1:-.*}"
CNT="${2:-170}"
timeout 2400 ncu --set full --clock-control none ${NCU_SOURCE:+--import-source on} -k "regex:$RE" -c "$CNT" -f -o gpurun_out/full \
    python scripts/step_once.py > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/full.ncu-rep --page raw --csv > gpurun_out/full_raw.csv 2> gpurun_out/full_raw.err
# gpurun brings back at most 64 MiB: keep the report only when it is small (NCU_SOURCE=1 with a narrow regex for source pages)
if [ "$(stat -c %s gpurun_out/full.ncu-rep 2>/dev/null || echo 0)" -gt 30000000 ]; then rm -f gpurun_out/full.ncu-rep; fi
ls -la gpurun_out/ | tail -5
