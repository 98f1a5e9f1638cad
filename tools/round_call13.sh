#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 bash tools/ab.sh base shw 2>&1 | grep -v "^\[" | tail -8 | cut -c1-200
export DSB_LIB_DIR=$PWD/desamba-so_b200/lib_shw
for ds in tiny mixed; do bash tools/gpu_check.sh $ds 2>&1 | grep PARITY; done
