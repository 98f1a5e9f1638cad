#!/bin/bash
# usage: tools/gpu_check.sh <dataset> [more datasets...]  -- CLI parity against the reference's golden SAM on the GPU box
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader | head -1
for ds in "$@"; do
  python tools/dataset.py "$ds" > gpurun_out/ds_$ds.json 2> gpurun_out/ds_$ds.log || { echo "dataset $ds failed"; tail -5 gpurun_out/ds_$ds.log; continue; }
  idx=$(python -c "import json;print(json.load(open('gpurun_out/ds_$ds.json'))['idx'])")
  dir=$(python -c "import json;print(json.load(open('gpurun_out/ds_$ds.json'))['dir'])")
  timeout 600 desamba-so_b200/bin/deSAMBA classify -g 1 -f SAM "$idx" "$dir/reads.fq" > gpurun_out/$ds.gpu.sam 2> gpurun_out/$ds.gpu.err
  echo "$ds rc=$?"; tail -3 gpurun_out/$ds.gpu.err
  if cmp -s "$dir/golden.sam" gpurun_out/$ds.gpu.sam; then echo "$ds PARITY OK ($(wc -l < $dir/golden.sam) lines)"; else
    echo "$ds PARITY DIFF"; diff <(cut -f1-6,12 "$dir/golden.sam") <(cut -f1-6,12 gpurun_out/$ds.gpu.sam) | head -20; cp "$dir/golden.sam" gpurun_out/$ds.golden.sam; fi
done
