#!/bin/bash
# tools/profile_kernel.sh TAG KERNEL_REGEX [SKIP]: one ncu --set full capture of a kernel of the short bench command
cd "$(dirname "$0")/.."
tag=$1; kern=$2; skip=${3:-3}
mkdir -p gpurun_out
SHORT="python bench.py --steps 2 --warmup 3 --no-fm-probe --no-cpu-baseline"
$SHORT > gpurun_out/plain_$tag.json 2> gpurun_out/plain_$tag.err &&
ncu --set full --clock-control none --import-source on -k regex:$kern -s $skip -c 1 -o gpurun_out/prof_$tag -f \
    $SHORT > gpurun_out/ncu_$tag.log 2>&1
tail -2 gpurun_out/ncu_$tag.log
