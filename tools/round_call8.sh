#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 bash tools/ab.sh base kbm1 kbp1 2>&1 | grep -v "^\[" | tail -12
