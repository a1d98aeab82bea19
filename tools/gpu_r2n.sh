#!/bin/bash
# round-2 GPU trip N (1 GPU): fused kernel with the second CTR key (launches that cross 2^32 keystream blocks)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused or whole_protocol or random_instance" > $O/r2n_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2n_pytest.log
timeout 120 python tools/kbench.py fused --n 1000000 --pairs 128 >> $O/r2n_kbench.txt 2>&1
timeout 120 python tools/kbench.py fused --n 125000 --pairs 1072 >> $O/r2n_kbench.txt 2>&1
cat $O/r2n_kbench.txt
exit 0
