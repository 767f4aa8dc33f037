#!/bin/bash
# long-sequence attention check: a few cells of the forward sweep (development helper)
timeout 300 python tools/sweep.py --precision f16 --out gpurun_out/sweep_cells.json --cells "2048x8,4096x4,16384x2" 2>&1 | grep '"N"' | cut -c1-260
