#!/bin/bash
# round-2 GPU trip H (1 GPU): seeded masks in the drop-in, balanced group rows
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_dropin.py tests/test_gpu_group.py -m gpu -x -q > $O/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 12 $O/r2h_pytest.log
P1GPU_TRACE=1 timeout 300 python tools/dropin_timing.py --n 1000000 --d 10 --providers 3 --seeded > $O/r2h_dropin_seeded_n1m_d10.json 2> $O/r2h_dropin_seeded.err; echo "dropin rc=$?"; cat $O/r2h_dropin_seeded_n1m_d10.json; grep "p1gpu trace" $O/r2h_dropin_seeded.err | tail -4
P1GPU_TRACE=1 timeout 300 python tools/dropin_timing.py --n 1000000 --d 20 --providers 4 --seeded > $O/r2h_dropin_seeded_n1m_d20.json 2> $O/r2h_dropin_seeded20.err; echo "dropin rc=$?"; cat $O/r2h_dropin_seeded_n1m_d20.json; grep "p1gpu trace" $O/r2h_dropin_seeded20.err | tail -5
exit 0
