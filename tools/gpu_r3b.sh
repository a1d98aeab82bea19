#!/bin/bash
# round 3 trip B: whole GPU suite (new: group double-reduce, sampled cross-pair words at cfg4/cfg5 size, no-alloc pack),
# smoke() with the extended kernel list, the default bench line (sampled cross-pair words of the timed cfg3 step, published block)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r3b_gpu_suite.log 2>&1; echo suite=$?; tail -3 gpurun_out/r3b_gpu_suite.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3b_smoke.log 2>&1; echo smoke=$?; tail -2 gpurun_out/r3b_smoke.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r3b_bench_cfg3.json 2> gpurun_out/r3b_bench_cfg3.err; echo bench=$?; tail -c 600 gpurun_out/r3b_bench_cfg3.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r3b_bench_cfg3.json').read().strip().splitlines()[-1])
print({k:l[k] for k in ('value','ms_per_step')}, l['e2e'], l['check'], l['roofline'])
PY
