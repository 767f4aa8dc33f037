"""Condense an `ncu --set full` report into the small per-launch CSV kept under profiles/ (run where ncu is installed).

    python tools/ncu_summary.py gpurun_out/r1e_gemm_mvp.ncu-rep profiles/r1e_gemm_mvp_ncu_full_summary.csv

Reads `ncu -i <rep> --page raw --csv` and keeps the columns DESIGN.md and bench.py's `roofline.traffic` quote: DRAM bytes,
duration, tensor-pipe activity, L2 sectors, shared-memory wavefronts, registers, smem, waves.  Row 2 of the output
carries the units, as in ncu's own export.
"""
from __future__ import annotations

import csv
import io
import subprocess
import sys

KEEP = [
    "Kernel Name", "Block Size", "Grid Size",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__time_duration.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
    "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum", "sm__cycles_elapsed.max",
    "sm__inst_executed.sum", "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_fma.sum",
    "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
]


def main() -> None:
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw[raw.index('"ID"'):])))
    head, units, body = rows[0], rows[1], rows[2:]
    cols = [head.index(k) for k in KEEP if k in head]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([head[c] for c in cols])
        w.writerow([units[c] for c in cols])
        for r in body:
            w.writerow([r[c][:70] if head[c] == "Kernel Name" else r[c] for c in cols])
    print(f"{out}: {len(body)} launches, {len(cols)} columns")


if __name__ == "__main__":
    main()
