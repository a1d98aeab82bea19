#!/bin/bash
# round-2 GPU trip M (8 GPUs): group and drop-in tests with 8 real devices (balanced row split, seeded masks)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_group.py tests/test_gpu_dropin.py -m gpu -q > $O/r2m_pytest_8gpu.log 2>&1; echo "pytest rc=$?"; tail -n 6 $O/r2m_pytest_8gpu.log
P1GPU_TRACE=1 timeout 300 python tools/dropin_timing.py --n 1000000 --d 10 --providers 3 --devices 0,1,2,3,4,5,6,7 --seeded > $O/r2m_dropin_seeded_8dev.json 2> $O/r2m_dropin_seeded_8dev.err; echo "dropin rc=$?"; cat $O/r2m_dropin_seeded_8dev.json
exit 0
