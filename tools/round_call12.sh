#!/bin/bash
cd "$(dirname "$0")/.."
export DSB_LIB_DIR=$PWD/desamba-so_b200/lib_base
timeout 900 bash tools/ab_env.sh "DSB_CARVEOUT=-1" "DSB_CARVEOUT=0" "DSB_CARVEOUT=25" 2>&1 | grep -v "^\[" | tail -12 | cut -c1-250
