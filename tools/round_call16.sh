#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 bash tools/ab.sh base mb6 mb7 2>&1 | grep -v "^\[" | tail -10 | cut -c1-200
