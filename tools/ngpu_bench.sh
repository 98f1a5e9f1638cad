#!/bin/bash
# tools/ngpu_bench.sh N: bench.py on N GPUs of one box the way the driver launches it (one rank per GPU under torchrun)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps ${2:-10} --warmup 3 \
    > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err || tail -8 gpurun_out/bench_${N}gpu.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_${N}gpu.json').read().strip().splitlines()[-1])
print('$N GPU: value %.0f e2e %.0f (%.2f ms) n_gpus %d kernels %.2f ms' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step'], d['n_gpus'], d['kernel_ms']['total']), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
