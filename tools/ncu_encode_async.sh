#!/bin/bash
# the cp.async encode kernel with and without digit planes, both ring widths, one launch each (single ncu capture)
set -e
cd "$(dirname "$0")/.."
python tools/kbench.py encode --n 1000000 --cols 500 --reps 1
python tools/kbench.py encode --n 1000000 --cols 500 --reps 1 --opt encode_planes=1
python tools/kbench.py encode --n 2000000 --cols 200 --bits 32 --precision 30 --reps 1 --opt encode_planes=1
