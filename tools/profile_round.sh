#!/bin/bash
# tools/profile_round.sh TAG: the round's measurement set on one B200 (run under gpurun).
#   gpurun_out/bench_TAG.json       full default bench line (fm probe + cpu baseline)
#   gpurun_out/launches_TAG.csv     every launch with its device time (ncu, cold-cache, serialised)
#   gpurun_out/prof_classify_TAG.ncu-rep   ncu --set full capture of one step's k_exist, k_islands, k_seed, k_classify
cd "$(dirname "$0")/.."
tag=${1:-rXX}
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err || { tail -5 gpurun_out/bench_$tag.err; exit 1; }
SHORT="python bench.py --steps 2 --warmup 3 --no-fm-probe --no-cpu-baseline"
$SHORT > gpurun_out/plain_$tag.json 2> gpurun_out/plain_$tag.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv \
    $SHORT > gpurun_out/ncu1_$tag.log 2>&1
$SHORT > gpurun_out/plain_$tag.json 2> gpurun_out/plain_$tag.err &&
ncu --set full --clock-control none --import-source on -k "regex:^k_islands|^k_seed|^k_classify|^k_exist$" -s 12 -c 4 -o gpurun_out/prof_classify_$tag -f \
    $SHORT > gpurun_out/ncu2_$tag.log 2>&1
tail -2 gpurun_out/ncu2_$tag.log
cat gpurun_out/bench_$tag.json | cut -c1-600
