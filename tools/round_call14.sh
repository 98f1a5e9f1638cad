#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
timeout 2000 python -m pytest tests -m gpu -x -q -s --durations=8 > gpurun_out/c14_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - t0 ))s"; tail -16 gpurun_out/c14_pytest.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "all done $(( $(date +%s) - t0 ))s"
