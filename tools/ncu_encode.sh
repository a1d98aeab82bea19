#!/bin/bash
# the encode pass for both ring widths, one launch each, for a single ncu --set full capture
set -e
cd "$(dirname "$0")/.."
python tools/kbench.py encode --n 1000000 --cols 500 --reps 1
python tools/kbench.py encode --n 2000000 --cols 200 --bits 32 --precision 30 --reps 1
