#!/bin/bash
# Round-2 parity additions on the GPU box: the reference itself (merge kernel object code, Python operators on cuda:0)
# and the bench scenes bit for bit against the oracle.  Logs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/summary_parity.txt
run() {
  echo "=== $1" | tee -a gpurun_out/summary_parity.txt
  timeout "$3" python -m pytest "$2" -m gpu -q -x --timeout "$3" > "gpurun_out/test_$1.log" 2>&1
  echo "exit $?" | tee -a gpurun_out/summary_parity.txt
  tail -n 15 "gpurun_out/test_$1.log" | tee -a gpurun_out/summary_parity.txt
}
nproc | tee -a gpurun_out/summary_parity.txt; free -g | head -2 | tee -a gpurun_out/summary_parity.txt
run reference_live tests/test_gpu_reference_live.py 900
run scale_parity tests/test_gpu_scale_parity.py 1500
cat gpurun_out/reference_live.json | tee -a gpurun_out/summary_parity.txt
