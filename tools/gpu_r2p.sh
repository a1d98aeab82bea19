#!/bin/bash
# round-2 GPU trip P (1 GPU): fused kernel with chunked work units
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused or whole_protocol or random_instance" > $O/r2p_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2p_pytest.log
for c in 1 2 4 8 16; do
timeout 120 python tools/kbench.py fused --n 1000000 --pairs 128 --opt fused_chunk=$c >> $O/r2p_kbench.txt 2>&1
done
timeout 120 python tools/kbench.py fused --n 125000 --pairs 1072 --opt fused_chunk=8 >> $O/r2p_kbench.txt 2>&1
timeout 120 python tools/kbench.py fused --n 125000 --pairs 1072 --opt fused_chunk=1 >> $O/r2p_kbench.txt 2>&1
cat $O/r2p_kbench.txt
exit 0
