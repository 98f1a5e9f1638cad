#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
timeout 900 python -m pytest tests -m gpu -x -q -k "not cfg2 and not sweep and not cfg3" > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - t0 ))s"; tail -3 gpurun_out/c2_pytest.log
python tools/e2e_trace.py 2> gpurun_out/c2_e2e_trace.txt; tail -12 gpurun_out/c2_e2e_trace.txt
python tools/e2e_sweep.py "" "DSB_BATCH_FIRST=4194304" "DSB_BATCH_TAPER=1" "DSB_BATCH_FIRST=4194304 DSB_BATCH_TAPER=1" "DSB_BATCH_BASES=13107200" "DSB_BATCH_BASES=8388608" "DSB_BATCH_BASES=8388608 DSB_BATCH_FIRST=4194304" "DSB_FORMAT_THREADS=12" 2> gpurun_out/c2_sweep.err | tee gpurun_out/c2_sweep.txt
echo "sweep done $(( $(date +%s) - t0 ))s"
python bench.py > gpurun_out/c2_bench_cfg2.json 2> gpurun_out/c2_bench_cfg2.err || tail -5 gpurun_out/c2_bench_cfg2.err
python -c "
import json; d=json.loads(open('gpurun_out/c2_bench_cfg2.json').read().strip().splitlines()[-1])
print('value %.0f e2e %.0f (%.2f ms) kernels' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step']), d['kernel_ms'])"
echo "all done $(( $(date +%s) - t0 ))s"
