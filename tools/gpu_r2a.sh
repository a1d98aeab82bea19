#!/bin/bash
# round-2 GPU trip A: new kernels' parity, then first numbers
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=12 > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -25 gpurun_out/r2a_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/r2a_bench_cfg3.json 2> gpurun_out/r2a_bench_cfg3.err; echo "bench rc=$?"
timeout 400 python bench.py --steps 2 --warmup 3 --aggregation masking-split --no-cpu-baseline --no-check --wire-sample 0 --ot-sample 0 > gpurun_out/r2a_bench_cfg3_split.json 2> gpurun_out/r2a_bench_cfg3_split.err; echo "split rc=$?"
timeout 400 python bench.py --steps 2 --warmup 3 --aggregation ot --ot-pairs 200 > gpurun_out/r2a_bench_cfg3_ot.json 2> gpurun_out/r2a_bench_cfg3_ot.err; echo "ot rc=$?"
timeout 300 python bench.py --workload gram --steps 5 --warmup 3 > gpurun_out/r2a_bench_gram.json 2> gpurun_out/r2a_bench_gram.err; echo "gram rc=$?"
for f in gpurun_out/r2a_bench_*.json; do echo "== $f"; python - "$f" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
except Exception as e:
    print("unreadable", e); sys.exit(0)
print({k:l.get(k) for k in ("value","ms_per_step","n_gpus","gpu_launches","value_with_wire")})
print("clocks",l.get("clocks")); print("check",l.get("check")); print("e2e",l.get("e2e")); print("int8",l.get("int8_peak_probe"))
print("roofline",l.get("roofline"))
for k,v in (l.get("kernels") or {}).items(): print(" ",k,{a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a!="lsu_bound"}, v.get("lsu_bound",{}).get("frac"))
print("use_ot",l.get("use_ot"), l.get("use_ot_sample")); print("cpu",l.get("cpu_baseline")); print("sustained",l.get("sustained"))
PY
done
tail -5 gpurun_out/*.err
