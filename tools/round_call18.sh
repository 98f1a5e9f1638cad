#!/bin/bash
cd "$(dirname "$0")/.."
R=$PWD/desamba-so_b200
python tools/e2e_sweep.py "DSB_LIB_DIR=$R/lib_mb6" "DSB_LIB_DIR=$R/lib_mb6s6" "DSB_LIB_DIR=$R/lib_mb5" | sed "s#$R/##"
DSB_WORKLOAD=cfg3s SWEEP_CALLS=5 python tools/e2e_sweep.py "DSB_LIB_DIR=$R/lib_mb6" "DSB_LIB_DIR=$R/lib_mb6s6" "DSB_LIB_DIR=$R/lib_mb5" | sed "s#$R/##"
timeout 600 bash tools/ab.sh mb6 mb6s6 mb5 2>&1 | grep -v "^\[" | grep classify | cut -c1-120
