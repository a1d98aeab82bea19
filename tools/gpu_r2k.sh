#!/bin/bash
# round-2 GPU trip K (1 GPU): the two bench arms exactly as the driver runs them
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
S=$(date +%s); timeout 900 python bench.py --impl reference --gpus 1 --steps 5 --warmup 3 > $O/r2k_bench_reference.json 2> $O/r2k_bench_reference.err; echo "reference rc=$? $(( $(date +%s) - S )) s"
S=$(date +%s); timeout 900 python bench.py --gpus 1 --steps 5 --warmup 3 > $O/r2k_bench_cfg3.json 2> $O/r2k_bench_cfg3.err; echo "bench rc=$? $(( $(date +%s) - S )) s"
for f in $O/r2k_bench_cfg3.json $O/r2k_bench_reference.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("impl","value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches","value_with_wire")})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e")); print("dropin",l.get("e2e_dropin"))
print("roofline",l.get("roofline"))
for k,v in (l.get("kernels") or {}).items(): print(" ",k,{a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("ms_per_step","avg_ms","frac","launches_per_step")}, (v.get("lsu_bound") or {}).get("frac"))
print("cpu",l.get("cpu_baseline")); print("wire",l.get("wire_format_sample")); print("ot",l.get("use_ot_sample"))
PY
done
tail -n 3 $O/r2k_*.err
exit 0
