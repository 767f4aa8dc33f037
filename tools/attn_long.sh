#!/bin/bash
# long-sequence attention A/B: sweep cells with and without the polynomial exponentials
for v in 0 1; do echo "AVD_ATTN_POLY=$v"; AVD_ATTN_POLY=$v timeout 300 python tools/sweep.py --out gpurun_out/sweep_poly$v.json --cells "2048x8,4096x4,16384x2" 2>&1 | grep '"N"' | cut -c1-260; done
