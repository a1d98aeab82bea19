#!/bin/bash
# one short launch of the kernels that make up the cfg3 step (mask expansion + the three party kernels),
# for a single `ncu --target-processes all --set full` capture
set -e
cd "$(dirname "$0")/.."
python tools/kbench.py ti --n 1000000 --pairs 33 --reps 1 --warmup 1
python tools/kbench.py party --n 1000000 --pairs 33 --reps 1 --warmup 1
