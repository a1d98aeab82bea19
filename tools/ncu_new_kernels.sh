#!/bin/bash
# one short launch of every kernel added after profiles/r1e (wire format, text ingestion) plus the
# encode pass, for a single `ncu --target-processes all` capture (see profiles/README in DESIGN.md §4)
set -e
cd "$(dirname "$0")/.."
python tools/kbench.py wire --n 1000000 --pairs 8 --reps 1 --warmup 1
python tools/kbench.py text --n 2000000 --reps 1
python tools/kbench.py encode --n 1000000 --cols 500 --reps 1
