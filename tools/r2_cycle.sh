#!/bin/bash
# Round-2 development cycle on the GPU box (through gpurun): GPU tests (all, not -x), A/B of the schedule switches on one
# DDIM step, the bench line with its extra legs.   usage: tools/r2_cycle.sh <tag> [pytest -k expression]
tag=${1:-r2x}
kexpr=${2:-}
o=gpurun_out
mkdir -p $o
if [ -n "$kexpr" ]; then
  timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -k "$kexpr" > $o/pytest_${tag}.log 2>&1
else
  timeout 1500 python -m pytest tests -m gpu -q --timeout 600 > $o/pytest_${tag}.log 2>&1
fi
tail -25 $o/pytest_${tag}.log
for opt in chains=1 chains=2 chains=4; do
  timeout 120 python tools/stepbench.py 8 audio --opt=$opt >> $o/stepbench_${tag}.log 2>&1
done
timeout 120 python tools/stepbench.py 64 video --opt=bx_wide=0 --kinds >> $o/stepbench_${tag}.log 2>&1
timeout 120 python tools/stepbench.py 64 video --opt=bx_wide=1 --kinds >> $o/stepbench_${tag}.log 2>&1
timeout 120 python tools/stepbench.py 64 video --opt=chains=2 >> $o/stepbench_${tag}.log 2>&1
cat $o/stepbench_${tag}.log
timeout 600 python bench.py > $o/${tag}_bench_v2a_b8.json 2> $o/${tag}_bench.err
python - <<PY
import json
try:
    r = json.loads(open("$o/${tag}_bench_v2a_b8.json").read().strip().splitlines()[-1])
    print("bench:", round(r["value"], 1), "clips/s  e2e", round(r["e2e"]["value"], 1), " gemm frac", round(r["roofline"]["frac"], 3), {k: v["us_per_launch"] for k, v in r["kernels"].items()})
    for k in ("strong_a2v_b64", "long_context", "torch_cuda_baseline", "cpu_baseline"):
        print(k, json.dumps(r.get(k))[:600])
except Exception as ex:
    print("bench failed", ex); print(open("$o/${tag}_bench.err").read()[-3000:])
PY
