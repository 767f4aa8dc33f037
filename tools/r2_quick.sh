#!/bin/bash
# Quick GPU check (through gpurun): selected GPU tests + one-step A/B lines.  usage: tools/r2_quick.sh <tag> <-k expr> [stepbench arg lines ...]
tag=$1; kexpr=$2; shift 2
o=gpurun_out; mkdir -p $o
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 -rP -k "$kexpr" > $o/pytest_${tag}.log 2>&1
grep -v "^$" $o/pytest_${tag}.log | tail -40
for line in "$@"; do
  timeout 180 python tools/stepbench.py $line >> $o/stepbench_${tag}.log 2>&1
done
cat $o/stepbench_${tag}.log
