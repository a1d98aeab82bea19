#!/bin/bash
# round-2 GPU trip G (8 GPUs): torchrun bench lines at N = 4 and 8 for the small-step workloads and the OT schedule,
# the group tests and the drop-in with 8 devices behind the C-ABI
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi -L > $O/r2g_gpus.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
SM="--no-cpu-baseline --no-dropin --wire-sample 0 --ot-sample 0"
port=29600
for N in 8 4; do
for W in cfg4 cfg5 gram; do
  port=$((port+1))
  timeout 240 $TR --nproc-per-node $N --master-port $port bench.py --gpus $N --workload $W --steps 5 --warmup 3 $SM > $O/r2g_bench_${W}_n$N.json 2> $O/r2g_bench_${W}_n$N.err; echo "$W n$N rc=$?"
done
port=$((port+1))
timeout 300 $TR --nproc-per-node $N --master-port $port bench.py --gpus $N --steps 2 --warmup 3 --aggregation ot --ot-pairs 1000 $SM > $O/r2g_bench_cfg3_ot_n$N.json 2> $O/r2g_bench_cfg3_ot_n$N.err; echo "ot n$N rc=$?"
done
P1GPU_TRACE=1 timeout 300 python tools/dropin_timing.py --n 1000000 --d 10 --providers 3 --devices 0,1,2,3,4,5,6,7 > $O/r2g_dropin_8dev.json 2> $O/r2g_dropin_8dev.err; echo "dropin 8dev rc=$?"
cat $O/r2g_dropin_8dev.json; grep "p1gpu trace" $O/r2g_dropin_8dev.err
timeout 600 python -m pytest tests/test_gpu_group.py tests/test_gpu_dropin.py -m gpu -x -q > $O/r2g_pytest_8gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 $O/r2g_pytest_8gpu.log
for f in $O/r2g_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches")})
print("sustained",{k:v for k,v in (l.get("sustained") or {}).items() if k!="clocks"})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e"))
PY
done
for f in $O/r2g_*.err; do echo "--- $f"; tail -n 3 "$f"; done
exit 0
