#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/c11_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - t0 ))s"; tail -3 gpurun_out/c11_pytest.log | cut -c1-300
python tools/cfg5_loop.py 100 cfg2 > gpurun_out/c11_cfg5.json 2> gpurun_out/c11_cfg5.err || tail -5 gpurun_out/c11_cfg5.err
cut -c1-900 gpurun_out/c11_cfg5.json
DSB_VERBOSE=1 python tools/cfg5_loop.py 3 cfg2 2>&1 >/dev/null | grep "meta_analysis:" | tail -2
echo "cfg5 done $(( $(date +%s) - t0 ))s"
bash tools/profile_round.sh r21 | cut -c1-400
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/c11_bench_ref.json 2> gpurun_out/c11_bench_ref.err; cut -c1-160 gpurun_out/c11_bench_ref.json
echo "all done $(( $(date +%s) - t0 ))s"
