#!/bin/bash
# round-2 GPU trip S (8 GPUs): the bench exactly as the driver's scaling run launches it at N = 8 (default flags), and cfg5 as
# BASELINE.json writes it (20 providers) at N = 8
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
S=$(date +%s); timeout 600 $TR --nproc-per-node 8 --master-port 29801 bench.py --gpus 8 --steps 5 --warmup 3 > $O/r2s_bench_cfg3_n8.json 2> $O/r2s_bench_cfg3_n8.err; echo "cfg3 n8 rc=$? $(( $(date +%s) - S )) s"
S=$(date +%s); timeout 600 $TR --nproc-per-node 8 --master-port 29802 bench.py --gpus 8 --impl reference --steps 5 --warmup 3 > $O/r2s_bench_reference_n8.json 2> $O/r2s_bench_reference_n8.err; echo "reference n8 rc=$? $(( $(date +%s) - S )) s"
S=$(date +%s); timeout 600 $TR --nproc-per-node 8 --master-port 29803 bench.py --gpus 8 --workload cfg5p20 --steps 2 --warmup 3 --no-cpu-baseline --no-dropin --wire-sample 0 --ot-sample 0 > $O/r2s_bench_cfg5p20_n8.json 2> $O/r2s_bench_cfg5p20_n8.err; echo "cfg5p20 n8 rc=$? $(( $(date +%s) - S )) s"
for f in $O/r2s_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("impl","value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches","value_with_wire")})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e")); print("batch", (l.get("config") or {}).get("batch_pairs"))
for k,v in (l.get("kernels") or {}).items(): print(" ",k,{a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("ms_per_step","avg_ms","frac","launches_per_step")})
PY
done
for f in $O/r2s_*.err; do echo "--- $f"; tail -n 3 "$f"; done
exit 0
