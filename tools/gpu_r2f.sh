#!/bin/bash
# round-2 GPU trip F (1 GPU): fused masking kernel at other CTA sizes
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
for t in 512 640 768 1024 384; do
timeout 120 python tools/kbench.py fused --n 1000000 --pairs 128 --opt fused_threads=$t >> $O/r2f_kbench2.txt 2>&1
done
cat $O/r2f_kbench2.txt
exit 0
