#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
export DSB_WORKLOAD=cfg3s SWEEP_CALLS=5
python tools/dataset.py cfg3s > /dev/null 2> gpurun_out/c3_ds.log; echo "dataset $(( $(date +%s) - t0 ))s"
python tools/e2e_sweep.py "" "DSB_BATCH_BASES=33554432" "DSB_BATCH_BASES=67108864" "DSB_SLOTS=6" "DSB_SLOTS=8" "DSB_BATCH_BASES=33554432 DSB_SLOTS=6" "DSB_FORMAT_THREADS=12" "DSB_BATCH_BASES=33554432 DSB_FORMAT_THREADS=12" "DSB_BATCH_BASES=8388608 DSB_SLOTS=8" | tee gpurun_out/c3_sweep_cfg3s.txt
echo "sweep done $(( $(date +%s) - t0 ))s"
DSB_VERBOSE=1 DSB_BATCH_BASES=33554432 python tools/e2e_sweep.py --one 2> gpurun_out/c3_trace_cfg3s_32M.txt; tail -30 gpurun_out/c3_trace_cfg3s_32M.txt | cut -c1-220
