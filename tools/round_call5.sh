#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/dataset.py cfg1 > gpurun_out/c5_ds.json 2> gpurun_out/c5_ds.log
idx=$(python -c "import json;print(json.load(open('gpurun_out/c5_ds.json'))['idx'])")
rd=$(python -c "import json;print(json.load(open('gpurun_out/c5_ds.json'))['reads'])")
CLI=desamba-so_b200/bin/deSAMBA
for e in "X=1" "DSB_SPLIT=0" "DSB_BATCH_BASES=16777216" "DSB_BATCH_READS=500"; do
  env $e $CLI classify -g 1 -f SAM $idx $rd > gpurun_out/c5_out.sam 2> gpurun_out/c5_err.txt; echo "$e rc=$? $(tail -1 gpurun_out/c5_err.txt | cut -c1-160)"
done
timeout 900 compute-sanitizer --tool memcheck --print-limit 5 $CLI classify -g 1 -f SAM $idx $rd > gpurun_out/c5_san.sam 2> gpurun_out/c5_san.txt; echo "sanitizer rc=$?"
grep -v "^=========     Host Frame\|^=========         in \|^=========                in " gpurun_out/c5_san.txt | head -60 | cut -c1-250
