#!/bin/bash
# tools/round_call1.sh: one gpurun call = GPU test suite, e2e host-pipeline timeline, launch-bound A/B, cfg3s parity + bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader | head -1; nproc; free -g | sed -n 2p
t0=$(date +%s)
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - t0 ))s"; tail -5 gpurun_out/c1_pytest.log
python tools/e2e_trace.py 2> gpurun_out/c1_e2e_trace.txt; tail -14 gpurun_out/c1_e2e_trace.txt
echo "trace done $(( $(date +%s) - t0 ))s"
timeout 600 bash tools/ab.sh base mb8 2>&1 | grep -v "^\[" | tail -8
echo "ab done $(( $(date +%s) - t0 ))s"
python bench.py > gpurun_out/c1_bench_cfg2.json 2> gpurun_out/c1_bench_cfg2.err || tail -5 gpurun_out/c1_bench_cfg2.err
cut -c1-400 gpurun_out/c1_bench_cfg2.json
echo "bench done $(( $(date +%s) - t0 ))s"
DSB_LONG_TESTS=1 timeout 1500 python -m pytest tests/test_gpu.py -m gpu -x -q -s -k cfg3 > gpurun_out/c1_cfg3s_parity.log 2>&1; echo "cfg3s rc=$?"; tail -4 gpurun_out/c1_cfg3s_parity.log
echo "cfg3s parity done $(( $(date +%s) - t0 ))s"
DSB_WORKLOAD=cfg3s python bench.py --steps 10 --warmup 3 > gpurun_out/c1_bench_cfg3s.json 2> gpurun_out/c1_bench_cfg3s.err || tail -5 gpurun_out/c1_bench_cfg3s.err
cut -c1-600 gpurun_out/c1_bench_cfg3s.json
echo "all done $(( $(date +%s) - t0 ))s"
