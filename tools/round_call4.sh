#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
timeout 900 python -m pytest tests -m gpu -x -q -k "not cfg2 and not sweep and not cfg3" > gpurun_out/c4_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - t0 ))s"; tail -3 gpurun_out/c4_pytest.log
python tools/e2e_trace.py 2> gpurun_out/c4_e2e_trace.txt; tail -13 gpurun_out/c4_e2e_trace.txt | cut -c1-200
python tools/e2e_sweep.py "" "DSB_BATCH_FIRST=4194304" "DSB_FORMAT_THREADS=8" | tee gpurun_out/c4_sweep_cfg2.txt
echo "sweep cfg2 done $(( $(date +%s) - t0 ))s"
DSB_WORKLOAD=cfg3s SWEEP_CALLS=5 python tools/e2e_sweep.py "" "DSB_SLOTS=6" "DSB_BATCH_BASES=16777216" | tee gpurun_out/c4_sweep_cfg3s.txt
echo "sweep cfg3s done $(( $(date +%s) - t0 ))s"
bash tools/profile_round.sh r19
echo "profile done $(( $(date +%s) - t0 ))s"
