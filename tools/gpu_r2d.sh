#!/bin/bash
# round-2 GPU trip D (1 GPU): whole GPU suite on the pooled allocator + new OT transpose, OT rate, bench lines, drop-in trace
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2d_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $O/r2d_pytest.log
timeout 120 python tools/kbench.py cot --n 1000000 --reps 5 >> $O/r2d_kbench.txt 2>&1
timeout 120 python tools/kbench.py cot --n 1000000 --reps 5 --bits 32 --precision 24 >> $O/r2d_kbench.txt 2>&1
cat $O/r2d_kbench.txt
P1GPU_TRACE=1 timeout 300 python tools/dropin_timing.py --n 1000000 --d 10 --providers 3 > $O/r2d_dropin_n1m_d10.json 2> $O/r2d_dropin.err; echo "dropin rc=$?"; cat $O/r2d_dropin_n1m_d10.json; grep "p1gpu trace" $O/r2d_dropin.err
P1GPU_TRACE=1 timeout 300 python tools/dropin_timing.py --n 1000000 --d 20 --providers 4 > $O/r2d_dropin_n1m_d20.json 2> $O/r2d_dropin20.err; echo "dropin20 rc=$?"; cat $O/r2d_dropin_n1m_d20.json; grep "p1gpu trace" $O/r2d_dropin20.err | tail -5
timeout 300 python tools/dropin_timing.py --n 100000 --d 10 --providers 3 --use-ot > $O/r2d_dropin_ot.json 2> $O/r2d_dropin_ot.err; echo "dropin ot rc=$?"; cat $O/r2d_dropin_ot.json
timeout 600 python bench.py --steps 3 --warmup 3 > $O/r2d_bench_cfg3.json 2> $O/r2d_bench_cfg3.err; echo "bench rc=$?"
timeout 600 python bench.py --steps 2 --warmup 3 --aggregation ot --ot-pairs 1000 --no-dropin > $O/r2d_bench_cfg3_ot.json 2> $O/r2d_bench_cfg3_ot.err; echo "ot rc=$?"
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 300 $NCU -k regex:k_cot -c 3 -o $O/r2d_cot python tools/kbench.py cot --n 1000000 --reps 1 --warmup 0 > $O/r2d_ncu_cot.log 2>&1
for f in $O/r2d_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches","value_with_wire")})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e")); print("dropin",l.get("e2e_dropin"))
print("roofline",l.get("roofline"))
for k,v in (l.get("kernels") or {}).items(): print(" ",k,{a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("ms_per_step","avg_ms","frac","launches_per_step")}, (v.get("lsu_bound") or {}).get("frac"))
print("use_ot",l.get("use_ot")); print("cpu",l.get("cpu_baseline"))
PY
done
for f in $O/r2d_*.err; do echo "--- $f"; tail -n 3 "$f"; done
exit 0
