#!/bin/bash
# round-2 GPU trip I (1 GPU): fused masking kernel with the fast unit loop
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused or whole_protocol or random_instance" > $O/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $O/r2i_pytest.log
for o in fused_fast=1 fused_fast=0; do
timeout 120 python tools/kbench.py fused --n 1000000 --pairs 128 --opt $o >> $O/r2i_kbench.txt 2>&1
done
timeout 120 python tools/kbench.py fused --n 1000000 --pairs 128 --opt fused_threads=640 >> $O/r2i_kbench.txt 2>&1
timeout 120 python tools/kbench.py fused --n 1000000 --pairs 128 --opt fused_threads=384 >> $O/r2i_kbench.txt 2>&1
cat $O/r2i_kbench.txt
exit 0
