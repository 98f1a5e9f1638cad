#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c19_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - t0 ))s"; tail -2 gpurun_out/c19_pytest.log | cut -c1-300
bash tools/profile_round.sh r23 | cut -c1-300
python tools/e2e_trace.py 2> gpurun_out/c19_e2e_trace.txt; grep -v "fill:" gpurun_out/c19_e2e_trace.txt | tail -17 | cut -c1-200
DSB_WORKLOAD=cfg3s python bench.py --steps 10 --warmup 3 --no-fm-probe > gpurun_out/c19_bench_cfg3s.json 2> gpurun_out/c19_bench_cfg3s.err
python -c "
import json; d=json.loads(open('gpurun_out/c19_bench_cfg3s.json').read().strip().splitlines()[-1])
print('cfg3s value %.0f e2e %.0f (%.2f ms)' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step']), d['kernel_ms'])"
python tools/cfg5_loop.py 60 cfg2 > gpurun_out/c19_cfg5.json 2>/dev/null; cut -c1-700 gpurun_out/c19_cfg5.json
echo "all done $(( $(date +%s) - t0 ))s"
