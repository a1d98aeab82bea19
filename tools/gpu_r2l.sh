#!/bin/bash
# round-2 GPU trip L (8 GPUs): cfg3 masking at N = 8, 4, 2 with the current kernels, the group reduce (P2P kernel vs NCCL)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
SM="--no-cpu-baseline --no-dropin --wire-sample 0 --ot-sample 0"
port=29700
for N in 8 4 2; do
  port=$((port+1))
  timeout 300 $TR --nproc-per-node $N --master-port $port bench.py --gpus $N --steps 5 --warmup 3 $SM > $O/r2l_bench_cfg3_n$N.json 2> $O/r2l_bench_cfg3_n$N.err; echo "cfg3 n$N rc=$?"
done
for D in 0,1 0,1,2,3 0,1,2,3,4,5,6,7; do for C in 200 500 1000; do timeout 120 python tools/kbench.py reduce --devices $D --cols $C --reps 200 >> $O/r2l_kbench_reduce.txt 2>&1; done; done
cat $O/r2l_kbench_reduce.txt
for f in $O/r2l_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches")})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e")); print("batch", l["config"].get("batch_pairs"))
PY
done
for f in $O/r2l_*.err; do echo "--- $f"; tail -n 3 "$f"; done
exit 0
