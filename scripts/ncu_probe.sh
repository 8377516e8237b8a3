#!/bin/bash
# launch list + `ncu --set full` (+ source) of the RLGR kernels matching a regex while scripts/rlgr_probe.py encodes the bench scene.
# usage: bash scripts/ncu_probe.sh <kernel-regex> [count] [probe args...] -> gpurun_out/launches_probe.csv, kp.ncu-rep, kp_raw.csv, kp_src.csv
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
RE="$1"; CNT="${2:-3}"; shift; shift
timeout 300 python scripts/rlgr_probe.py "$@" > gpurun_out/probe_plain.log 2>&1 || { echo "probe failed"; tail -5 gpurun_out/probe_plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:rlgr" --csv --log-file gpurun_out/launches_probe.csv python scripts/rlgr_probe.py "$@" > gpurun_out/ncu_probe.log 2>&1
python scripts/ncu_summary.py gpurun_out/launches_probe.csv > gpurun_out/launches_probe.txt; cat gpurun_out/launches_probe.txt
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$RE" -s 12 -c "$CNT" -f -o gpurun_out/kp python scripts/rlgr_probe.py "$@" > gpurun_out/ncu_kp.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/kp.ncu-rep --page raw --csv > gpurun_out/kp_raw.csv 2>/dev/null
ncu -i gpurun_out/kp.ncu-rep --page source --csv > gpurun_out/kp_src.csv 2>/dev/null
if [ "$(stat -c %s gpurun_out/kp.ncu-rep 2>/dev/null || echo 0)" -gt 30000000 ]; then rm -f gpurun_out/kp.ncu-rep; fi
ls -la gpurun_out | tail -6
