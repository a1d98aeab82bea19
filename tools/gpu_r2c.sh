#!/bin/bash
# round-2 GPU trip C (2 GPUs): the whole GPU suite with two real devices behind the group, the reduce
# micro-benchmark, torchrun bench lines at N = 1, 2 and the traced drop-in run
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi -L > $O/r2c_gpus.txt 2>&1
nvidia-smi topo -m >> $O/r2c_gpus.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2c_pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 $O/r2c_pytest_2gpu.log
for C in 200 500 1000; do timeout 120 python tools/kbench.py reduce --devices 0,1 --cols $C --reps 200 >> $O/r2c_kbench.txt 2>&1; done
cat $O/r2c_kbench.txt
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
SM="--no-cpu-baseline --no-dropin --wire-sample 0 --ot-sample 0"
for W in cfg4 cfg5 gram; do
  timeout 300 python bench.py --workload $W --steps 5 --warmup 3 $SM > $O/r2c_bench_${W}_n1.json 2> $O/r2c_bench_${W}_n1.err; echo "$W n1 rc=$?"
  timeout 300 $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2 --workload $W --steps 5 --warmup 3 $SM > $O/r2c_bench_${W}_n2.json 2> $O/r2c_bench_${W}_n2.err; echo "$W n2 rc=$?"
done
timeout 400 $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 $SM > $O/r2c_bench_cfg3_n2.json 2> $O/r2c_bench_cfg3_n2.err; echo "cfg3 n2 rc=$?"
timeout 400 $TR --nproc-per-node 2 --master-port 29513 bench.py --gpus 2 --steps 2 --warmup 3 --aggregation ot --ot-pairs 1000 $SM > $O/r2c_bench_cfg3_ot_n2.json 2> $O/r2c_bench_cfg3_ot_n2.err; echo "ot n2 rc=$?"
P1GPU_TRACE=1 timeout 300 python tools/dropin_timing.py --n 1000000 --d 10 --providers 3 > $O/r2c_dropin_1dev.json 2> $O/r2c_dropin_1dev.err; echo "dropin 1dev rc=$?"
P1GPU_TRACE=1 timeout 300 python tools/dropin_timing.py --n 1000000 --d 10 --providers 3 --devices 0,1 > $O/r2c_dropin_2dev.json 2> $O/r2c_dropin_2dev.err; echo "dropin 2dev rc=$?"
cat $O/r2c_dropin_1dev.json $O/r2c_dropin_2dev.json; grep "p1gpu trace" $O/r2c_dropin_1dev.err $O/r2c_dropin_2dev.err
for f in $O/r2c_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("value","ms_per_step","wall_ms_per_step","n_gpus","gpu_launches")})
print("sustained",{k:v for k,v in (l.get("sustained") or {}).items() if k!="clocks"})
print("clocks",l.get("clocks")); print("check",(l.get("check") or {}).get("ok")); print("e2e",l.get("e2e"))
for k,v in (l.get("kernels") or {}).items(): print(" ",k,{a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ("ms_per_step","avg_ms","frac","launches_per_step")})
print("use_ot",l.get("use_ot"))
PY
done
for f in $O/r2c_*.err; do echo "--- $f"; tail -n 4 "$f"; done
exit 0
