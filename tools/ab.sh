#!/bin/bash
# tools/ab.sh VARIANT...: same-box A/B of CUDA-layer variants built by tools/variant.sh.  For each variant: md5 of the
# SAM text read_classify() returns on the bench workload (must equal the first variant's), then two short bench runs.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for round in 1 2; do
  for v in "$@"; do
    export DSB_LIB_DIR=$PWD/desamba-so_b200/lib_$v
    if [ $round = 1 ]; then
      python - <<P
import sys, hashlib
sys.path.insert(0, "tools"); sys.path.insert(0, "desamba-so_b200")
import dataset, pydsb
m = dataset.build("cfg2", golden=False)
d = pydsb.Desamba(); d.load_index(m["idx"])
print("$v", hashlib.md5(d.read_classify(open(m["reads"], "rb").read(), thread_id=0, thread_num=1)).hexdigest(), flush=True)
P
    fi
    python bench.py --steps 5 --warmup 3 --no-fm-probe --no-cpu-baseline > gpurun_out/ab_${v}_$round.json 2> gpurun_out/ab_$v.err
    python -c "
import json; d=json.loads(open('gpurun_out/ab_${v}_$round.json').read().strip().splitlines()[-1])
print('$v', $round, 'exist %.2f classify %.2f total %.2f e2e %.2f' % (d['kernel_ms']['exist'], d['kernel_ms']['classify'], d['kernel_ms']['total'], d['e2e']['ms_per_step']), d['kernel_ms']['classify_stages'])"
  done
done
