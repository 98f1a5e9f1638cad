#!/bin/bash
# two GPUs of one box: peer-copied index, -g 2 in one process, bench.py under torchrun
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
t0=$(date +%s)
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 900 python -m pytest tests/test_gpu.py -m gpu -x -q -s -k "two_gpus or reblocking or cli_is_byte" > gpurun_out/c7_pytest.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - t0 ))s"; tail -3 gpurun_out/c7_pytest.log | cut -c1-300
python tools/dataset.py cfg3s > /dev/null 2> gpurun_out/c7_ds.log
idx=$(python -c "import json;print(json.load(open('/tmp/dsb_data/cfg3s/meta.json'))['idx'])")
for g in 1 2; do for i in 1 2; do DSB_VERBOSE=1 desamba-so_b200/bin/deSAMBA classify -g $g -f SAM_FULL $idx /tmp/dsb_data/cfg3s/reads.fq 2>&1 >/dev/null | grep "processed in\|index on GPU" | sed "s/^/-g $g: /"; done; done
DSB_INDEX_CLONE=0 DSB_VERBOSE=1 desamba-so_b200/bin/deSAMBA classify -g 2 -f SAM $idx /tmp/dsb_data/cfg3s/reads.fq 2>&1 >/dev/null | grep "index on GPU" | sed "s/^/no clone: /"
desamba-so_b200/bin/deSAMBA classify -g 1 -f SAM $idx /tmp/dsb_data/cfg3s/reads.fq 2>/dev/null | md5sum; desamba-so_b200/bin/deSAMBA classify -g 2 -f SAM $idx /tmp/dsb_data/cfg3s/reads.fq 2>/dev/null | md5sum
echo "cli done $(( $(date +%s) - t0 ))s"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c7_bench_2gpu.json 2> gpurun_out/c7_bench_2gpu.err || tail -5 gpurun_out/c7_bench_2gpu.err
python -c "
import json; d=json.loads(open('gpurun_out/c7_bench_2gpu.json').read().strip().splitlines()[-1])
print('2 GPU: value %.0f e2e %.0f (%.2f ms) n_gpus %d' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step'], d['n_gpus']), d['kernel_ms'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/c7_bench_ref_2gpu.json 2> gpurun_out/c7_bench_ref_2gpu.err; cut -c1-200 gpurun_out/c7_bench_ref_2gpu.json
echo "all done $(( $(date +%s) - t0 ))s"
