#!/bin/bash
cd "$(dirname "$0")/.."
timeout 1200 bash tools/ab.sh base hp1 hp3 h16 h16hp2 2>&1 | grep -v "^\[" | tail -20
