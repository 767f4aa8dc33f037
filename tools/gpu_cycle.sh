#!/bin/bash
# Development helper: GPU tests + kernel microbench + short bench on the GPU box (run through gpurun).
tag=${1:-run}
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_${tag}.log 2>&1; tail -6 gpurun_out/pytest_${tag}.log
timeout 120 python tools/kbench.py > gpurun_out/kbench_${tag}.log 2>&1; tail -9 gpurun_out/kbench_${tag}.log
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${tag}.json 2> gpurun_out/bench_${tag}.err
python - <<PY
import json
try:
    r = json.loads(open("gpurun_out/bench_${tag}.json").read().strip().splitlines()[-1])
    print("bench:", round(r["value"], 1), "clips/s  e2e", round(r["e2e"]["value"], 1), " gemm frac", round(r["roofline"]["frac"], 3), {k: v["us_per_launch"] for k, v in r["kernels"].items()})
except Exception as ex:
    print("bench failed", ex); print(open("gpurun_out/bench_${tag}.err").read()[-1500:])
PY
