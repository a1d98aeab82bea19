#!/bin/bash
# round-2 GPU trip E (1 GPU): whole-block one-digit Gram
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tcgen05 or local_block or planes" > $O/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $O/r2e_pytest.log
timeout 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "cfg4 or cfg5" > $O/r2e_pytest_full.log 2>&1; echo "pytest full rc=$?"; tail -n 4 $O/r2e_pytest_full.log
timeout 120 python tools/kbench.py gram --n 10000000 --cols 200 --target --bits 32 --precision 30 --reps 10 >> $O/r2e_kbench.txt 2>&1
timeout 120 python tools/kbench.py gram --n 10000000 --cols 200 --target --bits 32 --precision 30 --reps 10 --opt gram_tc_whole=0 >> $O/r2e_kbench.txt 2>&1
timeout 120 python tools/kbench.py gram --n 10000000 --cols 40 --target --bits 32 --precision 30 --reps 10 >> $O/r2e_kbench.txt 2>&1
timeout 120 python tools/kbench.py gram --n 10000000 --cols 40 --target --bits 32 --precision 30 --reps 10 --opt gram_tc_whole=0 >> $O/r2e_kbench.txt 2>&1
cat $O/r2e_kbench.txt
SM="--no-cpu-baseline --no-dropin --wire-sample 0 --ot-sample 0"
timeout 300 python bench.py --workload cfg4 --steps 5 --warmup 3 $SM > $O/r2e_bench_cfg4.json 2> $O/r2e_bench_cfg4.err; echo "cfg4 rc=$?"
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 300 $NCU -k regex:k_gram_whole -c 1 -o $O/r2e_gram_cfg4 python tools/kbench.py gram --n 10000000 --cols 200 --target --bits 32 --precision 30 --reps 1 --warmup 0 > $O/r2e_ncu_gramcfg4.log 2>&1
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r2e_bench_cfg4.json').read().strip().splitlines()[-1])
print({k:l.get(k) for k in ("value","ms_per_step","wall_ms_per_step")}, l['check']['ok'], l.get('sustained',{}).get('ms_per_step'))
for k,v in l['kernels'].items(): print(k, round(v['ms_per_step'],3), round(v.get('frac',0),3))
PY
tail -n 3 $O/r2e_*.err $O/r2e_ncu_gramcfg4.log
exit 0
