#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - t0 ))s"; tail -4 gpurun_out/c6_pytest.log | cut -c1-300
python tools/e2e_trace.py 2> gpurun_out/c6_e2e_trace.txt; tail -16 gpurun_out/c6_e2e_trace.txt | cut -c1-200
python bench.py > gpurun_out/c6_bench_cfg2.json 2> gpurun_out/c6_bench_cfg2.err || tail -5 gpurun_out/c6_bench_cfg2.err
python -c "
import json; d=json.loads(open('gpurun_out/c6_bench_cfg2.json').read().strip().splitlines()[-1])
print('value %.0f e2e %.0f (%.2f ms) kernels' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step']), d['kernel_ms'], d['cpu_baseline'])"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/c6_bench_ref.json 2> gpurun_out/c6_bench_ref.err; cut -c1-300 gpurun_out/c6_bench_ref.json
DSB_WORKLOAD=cfg3s python bench.py --steps 10 --warmup 3 --no-fm-probe > gpurun_out/c6_bench_cfg3s.json 2> gpurun_out/c6_bench_cfg3s.err
python -c "
import json; d=json.loads(open('gpurun_out/c6_bench_cfg3s.json').read().strip().splitlines()[-1])
print('cfg3s value %.0f e2e %.0f (%.2f ms) kernels' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step']), d['kernel_ms'], d['cpu_baseline'])"
# CLI throughput on files (mapped input)
idx=$(python -c "import json;print(json.load(open('/tmp/dsb_data/cfg3s/meta.json'))['idx'])")
for e in "X=1" "DSB_NO_MMAP=1"; do for i in 1 2; do env $e desamba-so_b200/bin/deSAMBA classify -g 1 -f SAM_FULL $idx /tmp/dsb_data/cfg3s/reads.fq 2>&1 >/dev/null | grep "processed in" | sed "s/^/$e /"; done; done
echo "all done $(( $(date +%s) - t0 ))s"
