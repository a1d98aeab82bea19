#!/bin/bash
# round-2 GPU trip T (2 GPUs): the bench with the driver's flags at N = 2, and the OT schedule at N = 2 with the current kernel
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
S=$(date +%s); timeout 300 $TR --nproc-per-node 2 --master-port 29811 bench.py --gpus 2 --steps 5 --warmup 3 > $O/r2t_bench_cfg3_n2.json 2> $O/r2t_bench_cfg3_n2.err; echo "cfg3 n2 rc=$? $(( $(date +%s) - S )) s"
S=$(date +%s); timeout 300 $TR --nproc-per-node 2 --master-port 29812 bench.py --gpus 2 --steps 2 --warmup 3 --aggregation ot --ot-pairs 1000 --no-cpu-baseline --no-dropin --no-e2e > $O/r2t_bench_cfg3_ot_n2.json 2> $O/r2t_bench_cfg3_ot_n2.err; echo "ot n2 rc=$? $(( $(date +%s) - S )) s"
for f in $O/r2t_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("impl","value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches","value_with_wire")})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e")); print("batch", (l.get("config") or {}).get("batch_pairs"))
PY
done
for f in $O/r2t_*.err; do echo "--- $f"; tail -n 3 "$f"; done
exit 0
