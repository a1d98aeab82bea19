#!/bin/bash
# round 3 trip H: the default line and cfg4 with the seeded numpy instance generator
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/r3h_bench_cfg3.json 2> gpurun_out/r3h_bench_cfg3.err; echo bench=$?; tail -c 600 gpurun_out/r3h_bench_cfg3.err
python - <<'PY'
import json
l=json.loads(open("gpurun_out/r3h_bench_cfg3.json").read().strip().splitlines()[-1])
print({k:l[k] for k in ("value","ms_per_step")}, l["e2e"]["ms_per_step"], l["config"]["data_seed"][:60])
print("CHECK", l["check"]["ok"], l["check"]["sampled_cross_pair_words_match"], "OT", l["ot_workload"]["cot_per_s"], l["ot_workload"]["check"]["ok"], "CPU", l["cpu_baseline"]["gpu_shares_match_sample"], l["cpu_baseline"]["value"])
print("GRAM", l["gram_p1"]["back_to_back"]["frac_of_nominal_dense_int8"], "dropin", l["e2e_dropin"]["speedup_warm"], l["clocks"])
PY
python bench.py --workload cfg4 --steps 5 --warmup 3 --no-cpu-baseline --no-dropin > gpurun_out/r3h_bench_cfg4.json 2> gpurun_out/r3h_bench_cfg4.err; echo cfg4=$?; tail -c 300 gpurun_out/r3h_bench_cfg4.err
python - <<'PY'
import json
l=json.loads(open("gpurun_out/r3h_bench_cfg4.json").read().strip().splitlines()[-1])
print({k:l[k] for k in ("value","ms_per_step")}, l["check"]["ok"], l["roofline"])
PY
