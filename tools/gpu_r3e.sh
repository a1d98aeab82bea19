#!/bin/bash
# round 3 trip E (8 GPUs): group tests on real peers (concurrent bring-up, reduce twice), the drop-in behind the reference's
# entry points with 8 devices (cold / warm), the default bench line at N = 8 (sampled cross-pair words, ot_workload, gram_p1)
mkdir -p gpurun_out; O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_group.py -m gpu -x -q > $O/r3e_pytest_group_8gpu.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/r3e_pytest_group_8gpu.log
P1GPU_TRACE=1 timeout 300 python tools/dropin_timing.py --n 1000000 --d 10 --providers 3 --devices 0,1,2,3,4,5,6,7 > $O/r3e_dropin_8dev.json 2> $O/r3e_dropin_8dev.err; echo "dropin 8dev rc=$?"
cat $O/r3e_dropin_8dev.json
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 5 --warmup 3 > $O/r3e_bench_cfg3_n8.json 2> $O/r3e_bench_cfg3_n8.err; echo "bench rc=$?"; tail -c 600 $O/r3e_bench_cfg3_n8.err
python - <<'PY'
import json
l=json.loads(open("gpurun_out/r3e_bench_cfg3_n8.json").read().strip().splitlines()[-1])
print({k:l[k] for k in ("value","ms_per_step","n_gpus")}, l["e2e"]["ms_per_step"])
print("CHECK", {k:v for k,v in l["check"].items() if k not in ("what","sampled_cross_pair_q")})
o=l["ot_workload"]; print("OT", o["inner_products"], o["ms"], o["cot_per_s"], o["session_setup_ms"], o["workload_seconds_extrapolated"], o["check"]["ok"])
print("GRAM", l["gram_p1"]["back_to_back"], l["gram_p1"]["ms_back_to_back"])
PY
