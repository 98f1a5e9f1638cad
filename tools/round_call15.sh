#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
N=${1:-4}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/c15_bench_${N}gpu.json 2> gpurun_out/c15_bench_${N}gpu.err || tail -8 gpurun_out/c15_bench_${N}gpu.err
python -c "
import json; d=json.loads(open('gpurun_out/c15_bench_${N}gpu.json').read().strip().splitlines()[-1])
print('$N GPU: value %.0f e2e %.0f (%.2f ms) n_gpus %d' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step'], d['n_gpus']), d['kernel_ms']['total'], d['clocks'])"
echo "bench done $(( $(date +%s) - t0 ))s"
idx=$(python -c "import json;print(json.load(open('/tmp/dsb_data/cfg2/meta.json'))['idx'])")
cat /tmp/dsb_data/cfg2/reads.fq /tmp/dsb_data/cfg2@1/reads.fq /tmp/dsb_data/cfg2@2/reads.fq /tmp/dsb_data/cfg2@3/reads.fq > /tmp/big.fq
for g in 1 $N; do desamba-so_b200/bin/deSAMBA classify -g $g -f SAM_FULL $idx /tmp/big.fq 2>&1 >/dev/null | grep "processed in" | sed "s/^/-g $g: /"; done
desamba-so_b200/bin/deSAMBA classify -g 1 -f SAM $idx /tmp/big.fq 2>/dev/null | md5sum; desamba-so_b200/bin/deSAMBA classify -g $N -f SAM $idx /tmp/big.fq 2>/dev/null | md5sum
echo "all done $(( $(date +%s) - t0 ))s"
