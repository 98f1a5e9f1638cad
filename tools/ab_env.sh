#!/bin/bash
# tools/ab_env.sh "VAR=val ..." "VAR=val ..." : same-box A/B of run-time settings of the built library (see ab.sh)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
i=0
for round in 1 2; do
  i=0
  for v in "$@"; do
    i=$((i+1))
    if [ $round = 1 ]; then
      env $v python - <<P
import sys, hashlib
sys.path.insert(0, "tools"); sys.path.insert(0, "desamba-so_b200")
import dataset, pydsb
m = dataset.build("cfg2", golden=False)
d = pydsb.Desamba(); d.load_index(m["idx"])
print("$v", hashlib.md5(d.read_classify(open(m["reads"], "rb").read(), thread_id=0, thread_num=1)).hexdigest(), flush=True)
P
    fi
    env $v python bench.py --steps 5 --warmup 3 --no-fm-probe --no-cpu-baseline > gpurun_out/abe_${i}_$round.json 2> gpurun_out/abe_$i.err || tail -5 gpurun_out/abe_$i.err
    python -c "
import json; d=json.loads(open('gpurun_out/abe_${i}_$round.json').read().strip().splitlines()[-1])
print('$v', $round, 'exist %.2f classify %.2f total %.2f e2e %.2f' % (d['kernel_ms']['exist'], d['kernel_ms']['classify'], d['kernel_ms']['total'], d['e2e']['ms_per_step']), d['kernel_ms']['classify_stages'])"
  done
done
