#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "not cfg2 and not sweep and not cfg3 and not beyond and not larger" > gpurun_out/c21_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/c21_pytest.log
python bench.py > gpurun_out/bench_r24.json 2> gpurun_out/bench_r24.err || tail -5 gpurun_out/bench_r24.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_r24.json').read().strip().splitlines()[-1])
print('value %.0f e2e %.0f (%.2f ms)' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step']), d['kernel_ms'], d['cpu_baseline']['value'], d['roofline']['traffic'])"
