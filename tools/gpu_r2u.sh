#!/bin/bash
# round-2 GPU trip U (1 GPU): cfg2 and cfg5 with 20 providers on the final kernels
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
SM="--no-cpu-baseline --no-dropin --wire-sample 0 --ot-sample 0"
timeout 200 python bench.py --workload cfg2 --steps 5 --warmup 3 $SM > $O/r2u_bench_cfg2.json 2> $O/r2u_bench_cfg2.err; echo "cfg2 rc=$?"
timeout 300 python bench.py --workload cfg5p20 --steps 2 --warmup 3 $SM > $O/r2u_bench_cfg5p20.json 2> $O/r2u_bench_cfg5p20.err; echo "cfg5p20 rc=$?"
for f in $O/r2u_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k:l.get(k) for k in ("value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches")})
print("sustained",{k:v for k,v in (l.get("sustained") or {}).items() if k!="clocks"}); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e")); print("lsu", l["roofline"].get("lsu_bound",{}).get("frac"))
PY
done
exit 0
