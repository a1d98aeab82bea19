#!/bin/bash
# round-2 GPU trip B: group / drop-in tests, new OT kernel, fused-kernel tuning, ncu captures
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_group.py tests/test_gpu_dropin.py -m gpu -x -q > $O/r2b_pytest_group.log 2>&1; echo "pytest group rc=$?"; tail -5 $O/r2b_pytest_group.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "cot or fused or whole or shards" > $O/r2b_pytest_parity.log 2>&1; echo "pytest parity rc=$?"; tail -3 $O/r2b_pytest_parity.log
timeout 300 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "use_ot" > $O/r2b_pytest_ot.log 2>&1; echo "pytest ot rc=$?"; tail -3 $O/r2b_pytest_ot.log
for B in 33 64 128 256; do timeout 120 python tools/kbench.py fused --n 1000000 --pairs $B --reps 5 >> $O/r2b_kbench.txt 2>&1; done
timeout 120 python tools/kbench.py fused --n 1000000 --pairs 128 --reps 5 --opt fused_threads=384 >> $O/r2b_kbench.txt 2>&1
timeout 120 python tools/kbench.py ti --n 1000000 --pairs 128 --reps 5 >> $O/r2b_kbench.txt 2>&1
timeout 120 python tools/kbench.py cot --n 1000000 --reps 5 >> $O/r2b_kbench.txt 2>&1
timeout 120 python tools/kbench.py cot --n 1000000 --reps 5 --bits 32 --precision 24 >> $O/r2b_kbench.txt 2>&1
timeout 120 python tools/kbench.py probe >> $O/r2b_kbench.txt 2>&1
timeout 120 python tools/kbench.py gram --n 1000000 --cols 500 --target --reps 10 >> $O/r2b_kbench.txt 2>&1
timeout 120 python tools/kbench.py gram --n 1000000 --cols 1000 --target --reps 10 >> $O/r2b_kbench.txt 2>&1
cat $O/r2b_kbench.txt
timeout 500 python bench.py --steps 3 --warmup 3 > $O/r2b_bench_cfg3.json 2> $O/r2b_bench_cfg3.err; echo "bench rc=$?"
timeout 400 python bench.py --steps 2 --warmup 3 --aggregation ot --ot-pairs 300 > $O/r2b_bench_cfg3_ot.json 2> $O/r2b_bench_cfg3_ot.err; echo "ot rc=$?"
timeout 300 python tools/dropin_timing.py --n 1000000 --d 10 --providers 3 > $O/r2b_dropin_n1m_d10.json 2> $O/r2b_dropin.err; echo "dropin rc=$?"; cat $O/r2b_dropin_n1m_d10.json
# ---- ncu --set full captures (one GPU, after the plain runs above exited)
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 300 $NCU -k regex:k_masking_fused -c 1 -o $O/r2b_fused python tools/kbench.py fused --n 1000000 --pairs 33 --reps 1 --warmup 0 > $O/r2b_ncu_fused.log 2>&1
timeout 300 $NCU -k regex:k_cot -c 3 -o $O/r2b_cot python tools/kbench.py cot --n 1000000 --reps 1 --warmup 0 > $O/r2b_ncu_cot.log 2>&1
timeout 300 $NCU -k regex:k_gram_tc -c 1 -o $O/r2b_gram_d500 python tools/kbench.py gram --n 1000000 --cols 500 --target --reps 1 --warmup 0 > $O/r2b_ncu_gram500.log 2>&1
timeout 300 $NCU -k regex:k_gram_tc -c 1 -o $O/r2b_gram_d1000 python tools/kbench.py gram --n 1000000 --cols 1000 --target --reps 1 --warmup 0 > $O/r2b_ncu_gram1000.log 2>&1
timeout 300 $NCU -k regex:k_gram_tc -c 1 -o $O/r2b_gram_cfg4 python tools/kbench.py gram --n 10000000 --cols 200 --target --bits 32 --precision 30 --reps 1 --warmup 0 > $O/r2b_ncu_gramcfg4.log 2>&1
ls -la $O/*.ncu-rep
for f in $O/r2b_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("value","ms_per_step","n_gpus","gpu_launches","value_with_wire")})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e"))
for k,v in (l.get("kernels") or {}).items(): print(" ",k,{a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("ms_per_step","avg_ms","frac","launches_per_step")}, v.get("lsu_bound",{}).get("frac"))
print("use_ot",l.get("use_ot"))
PY
done
for f in $O/r2b_*.err; do echo "--- $f"; tail -n 3 "$f"; done
exit 0
