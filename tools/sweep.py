"""MMDT forward roofline sweep on one B200 (BASELINE.json configs[4]; development/measurement tool).

    python tools/sweep.py [--out gpurun_out/sweep.json] [--max-tokens 131072] [--cells 2048x8,4096x4]

For joint sequence length N in {512 .. 16384} x batch B in {1 .. 64} (cells above --max-tokens rows are skipped) it
runs the mvp.yaml MMDiT (d = 512, L = 8, H = 8, MLP 2048, random init) forward in --precision mode (bf16 / f16) through
avd_mmdt_blocks on unit-rms activations, times it with CUDA events (warm-up, then the median of `iters` runs), and
times the GEMM family and the attention family separately by replaying the same call with the other kernels
dropped (avd_core_workspace.profile_skip_kinds).  FLOPs are SURVEY.md §8(d)'s algorithmic figures: GEMM 24 d^2 per token-layer,
attention 4 N d per token-layer.  Peak = MEASURED_PEAKS.json sustained bf16.  (Correctness of the same call is the
job of tests/test_gpu_model.py; this tool only measures.)
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import statistics
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from multimodal_diffusion_b200 import _lib, packing  # noqa: E402
from multimodal_diffusion_b200.avdiff.models.mmdt import MMDiT  # noqa: E402

KIND_GEMM, KIND_ATTN_TC = 2, 3


def time_call(fn, iters: int) -> float:
    """median microseconds of fn() over `iters` event-timed runs on the current stream."""
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/sweep.json")
    ap.add_argument("--max-tokens", type=int, default=131072)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--cells", default="", help="comma-separated NxB cells to run instead of the whole grid")
    args = ap.parse_args()

    dev = torch.device("cuda:0")
    lib = _lib.load()
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    peak = float(peaks.get("bf16_tflops_sustained", 1379.5))
    d, L, H, dm = 512, 8, 8, 2048
    torch.manual_seed(102)
    core = MMDiT(d_model=d, n_layers=L, n_heads=H, mlp_ratio=4.0, dropout=0.0).eval().to(dev)
    prec = _lib.precision_code(args.precision)
    pc = packing.PackedCore(core.state_dict(), L, H, prec, dev)
    n_kinds = lib.avd_prof_num_kinds()
    all_kinds = (1 << n_kinds) - 1

    rows = []
    grid = [(N, B) for N in (512, 1024, 2048, 4096, 8192, 16384) for B in (1, 2, 4, 8, 16, 32, 64)]
    if args.cells:
        grid = [tuple(int(v) for v in c.split("x")) for c in args.cells.split(",")]
    for N, B in grid:
        if True:
            M = B * N
            if M > args.max_tokens:
                continue
            x = torch.randn(M, d, device=dev)
            xa, ssq = packing.rows_prepare(x, prec)
            ws, bufs = packing.core_workspace(M, d, dm, prec, dev)

            def fwd():
                _lib.check(lib.avd_mmdt_blocks(C.byref(pc.struct), x.data_ptr(), xa.data_ptr(), ssq.data_ptr(),
                                               C.byref(ws), B, N, _lib.stream_ptr()))

            us_all = time_call(fwd, args.iters)
            ws.profile_skip_kinds = all_kinds & ~(1 << KIND_GEMM)      # scoped to this workspace's calls
            us_gemm = time_call(fwd, args.iters)
            ws.profile_skip_kinds = all_kinds & ~(1 << KIND_ATTN_TC)
            us_attn = time_call(fwd, args.iters)
            ws.profile_skip_kinds = 0
            f_gemm = L * M * 24.0 * d * d
            f_attn = L * M * 4.0 * N * d
            row = {
                "N": N, "B": B, "tokens": M,
                "fwd_us": round(us_all, 1), "gemm_us": round(us_gemm, 1), "attn_us": round(us_attn, 1),
                "fwd_tflops": round((f_gemm + f_attn) / us_all * 1e-6, 1),
                "gemm_tflops": round(f_gemm / us_gemm * 1e-6, 1),
                "attn_tflops": round(f_attn / us_attn * 1e-6, 1),
                "attn_flop_share": round(f_attn / (f_gemm + f_attn), 3),
            }
            row["fwd_frac"] = round(row["fwd_tflops"] / peak, 3)
            row["gemm_frac"] = round(row["gemm_tflops"] / peak, 3)
            row["attn_frac"] = round(row["attn_tflops"] / peak, 3)
            rows.append(row)
            print(json.dumps(row), flush=True)
            del x, xa, ssq, ws, bufs
    out = {"what": f"MMDT forward (mvp.yaml core, {args.precision} mode) on 1 x B200: CUDA-event medians; gemm/attn = same "
                   "call with the other kernel families dropped", "peak_tflops": peak,
           "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained", "precision": args.precision, "rows": rows}
    Path(args.out).parent.mkdir(parents=True, exist_ok=True)
    Path(args.out).write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
