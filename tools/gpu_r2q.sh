#!/bin/bash
# round-2 GPU trip Q (1 GPU): final state - whole GPU suite, smoke, both bench arms as the driver runs them, ncu of the fused kernel
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2q_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2q_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2q_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 $O/r2q_smoke.log
S=$(date +%s); timeout 900 python bench.py --gpus 1 --steps 5 --warmup 3 > $O/r2q_bench_cfg3.json 2> $O/r2q_bench_cfg3.err; echo "bench rc=$? $(( $(date +%s) - S )) s"
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 300 $NCU -k regex:k_masking_fused -c 1 -o $O/r2q_fused python tools/kbench.py fused --n 1000000 --pairs 33 --reps 1 --warmup 0 > $O/r2q_ncu_fused.log 2>&1
for f in $O/r2q_bench_cfg3.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("impl","value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches","value_with_wire")})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e"))
print("dropin",{k:v for k,v in (l.get("e2e_dropin") or {}).items() if k in ("gpu_host_warm_s","reference_phase1_c_s","speedup_warm","shares_identical")})
print("lsu",l["roofline"]["lsu_bound"]["frac"], "hbm frac", l["roofline"]["frac"])
for k,v in (l.get("kernels") or {}).items(): print(" ",k,{a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("ms_per_step","avg_ms","frac","launches_per_step")})
PY
done
tail -n 3 $O/r2q_*.err
exit 0
