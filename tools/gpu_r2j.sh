#!/bin/bash
# round-2 GPU trip J (1 GPU): the whole GPU suite, the default bench line (as the driver runs it), the reference arm,
# ncu capture of the fused kernel with the fast loop, launch list of the bench command
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2j_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $O/r2j_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2j_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $O/r2j_smoke.log
/usr/bin/time -v timeout 900 python bench.py > $O/r2j_bench_cfg3.json 2> $O/r2j_bench_cfg3.err; echo "bench rc=$?"; grep "Elapsed" $O/r2j_bench_cfg3.err
timeout 600 python bench.py --steps 2 --warmup 3 --aggregation ot --ot-pairs 1000 --no-dropin > $O/r2j_bench_cfg3_ot1000.json 2> $O/r2j_bench_cfg3_ot.err; echo "ot rc=$?"
SM="--no-cpu-baseline --no-dropin --wire-sample 0 --ot-sample 0"
for W in cfg2 cfg4 cfg5 gram cfg4p5 cfg5p20; do
  timeout 400 python bench.py --workload $W --steps 3 --warmup 3 $SM > $O/r2j_bench_${W}.json 2> $O/r2j_bench_${W}.err; echo "$W rc=$?"
done
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 300 $NCU -k regex:k_masking_fused -c 1 -o $O/r2j_fused python tools/kbench.py fused --n 1000000 --pairs 33 --reps 1 --warmup 0 > $O/r2j_ncu_fused.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/r2j_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-dropin --wire-sample 0 --ot-sample 0 --no-e2e > $O/r2j_ncu_bench.log 2>&1; echo "ncu list rc=$?"
for f in $O/r2j_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches","value_with_wire")})
print("sustained",{k:v for k,v in (l.get("sustained") or {}).items() if k!="clocks"})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e")); print("dropin",l.get("e2e_dropin"))
print("roofline",l.get("roofline"))
for k,v in (l.get("kernels") or {}).items(): print(" ",k,{a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("ms_per_step","avg_ms","frac","launches_per_step")}, (v.get("lsu_bound") or {}).get("frac"))
print("cpu",l.get("cpu_baseline"))
PY
done
for f in $O/r2j_*.err; do echo "--- $f"; tail -n 3 "$f"; done
exit 0
