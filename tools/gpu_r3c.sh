#!/bin/bash
# round 3 trip C: suite again (pool scrub, staged peer tables), default bench line with ot_workload + gram_p1 blocks,
# reference arm at the driver's flags (must say d=10)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r3c_gpu_suite.log 2>&1; echo suite=$?; tail -3 gpurun_out/r3c_gpu_suite.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r3c_bench_cfg3.json 2> gpurun_out/r3c_bench_cfg3.err; echo bench=$?; tail -c 1500 gpurun_out/r3c_bench_cfg3.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/r3c_bench_cfg3.json').read().strip().splitlines()[-1])
print({k:l[k] for k in ('value','ms_per_step')}, l['e2e'])
print('OT', l['ot_workload'])
print('GRAM', l['gram_p1'])
print('PUB', l['published'])
PY
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r3c_bench_reference.json 2> gpurun_out/r3c_bench_reference.err; echo ref=$?; cut -c1-900 gpurun_out/r3c_bench_reference.json
