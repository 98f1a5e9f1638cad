#!/bin/bash
# tools/ab_blocks.sh: same-box A/B of the persistent grids' width (DSB_BLOCKS_PER_SM for k_islands / k_classify,
# DSB_SEED_BLOCKS_PER_SM for k_seed).  A narrower k_classify leaves registers for the next batch's k_exist / k_seed on the
# same SM: the kernels of batches in flight can then overlap instead of queueing behind a full device.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for round in 1 2; do
  for cfg in "$@"; do
    b=${cfg%%:*}; sd=${cfg##*:}
    DSB_BLOCKS_PER_SM=$b DSB_SEED_BLOCKS_PER_SM=$sd python bench.py --steps 5 --warmup 3 --no-fm-probe --no-cpu-baseline > gpurun_out/abb_${b}_${sd}_$round.json 2> gpurun_out/abb_${b}_${sd}.err
    python -c "
import json; d=json.loads(open('gpurun_out/abb_${b}_${sd}_$round.json').read().strip().splitlines()[-1])
print('classify $b/SM seed $sd/SM round $round: value %.1f ms, h2d..d2h %.1f ms, e2e %.1f ms' % (d['ms_per_step'], d['h2d_kernels_d2h']['ms_per_step'], d['e2e']['ms_per_step']), d['kernel_ms']['classify_stages'])"
  done
done
