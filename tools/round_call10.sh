#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/c10_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - t0 ))s"; tail -4 gpurun_out/c10_pytest.log | cut -c1-400
timeout 900 bash tools/ab.sh nodp regdp 2>&1 | grep -v "^\[" | tail -8
echo "all done $(( $(date +%s) - t0 ))s"
