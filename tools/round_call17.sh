#!/bin/bash
cd "$(dirname "$0")/.."
R=$PWD/desamba-so_b200
python tools/e2e_sweep.py "DSB_LIB_DIR=$R/lib_base" "DSB_LIB_DIR=$R/lib_mb6" "DSB_LIB_DIR=$R/lib_mb7" | sed "s#$R/##"
DSB_WORKLOAD=cfg3s SWEEP_CALLS=5 python tools/e2e_sweep.py "DSB_LIB_DIR=$R/lib_base" "DSB_LIB_DIR=$R/lib_mb6" "DSB_LIB_DIR=$R/lib_mb7" | sed "s#$R/##"
