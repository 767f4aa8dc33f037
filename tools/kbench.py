"""Kernel micro-benchmarks on one B200 (development tool, not part of the product path).

    python tools/kbench.py [--trace] [--reps 40]

For every GEMM / attention shape of one mvp-shape DDIM step (V2A, B = 8 -> M = 2128 rows) it captures a CUDA
graph of `reps` back-to-back launches, replays it and prints microseconds per launch (warm L2) and the
achieved TFLOP/s.  --trace additionally prints the in-kernel phase stamps of the GEMM (avd_debug_trace).
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from multimodal_diffusion_b200 import _lib, packing  # noqa: E402


def time_graph(fn, reps: int, iters: int = 20) -> float:
    """microseconds per call of fn inside a CUDA graph of `reps` calls."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (iters * reps)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=40)
    ap.add_argument("--trace", action="store_true")
    ap.add_argument("--M", type=int, default=2128)
    ap.add_argument("--once", action="store_true", help="run one block's kernels 3 times eagerly and exit (ncu target)")
    args = ap.parse_args()
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    M, d, dm = args.M, 512, 2048
    P = _lib.AVD_PREC_BF16

    def mk(rows, cols, S=1):
        return (torch.randn(S, rows, cols, device=dev) * 0.05).to(torch.bfloat16).contiguous()

    x = torch.randn(M, d, device=dev)
    ssq = torch.rand(M, d // 64, device=dev) + 0.5
    shapes = []
    # name, M, N, K, kwargs
    A512, A2048 = mk(M, d), mk(M, dm)
    Wqkv, Wo, W1, W2 = mk(3 * d, d), mk(d, d), mk(dm, d), mk(d, dm)
    bq, bo, b1, b2 = (torch.randn(n, device=dev) for n in (3 * d, d, dm, d))
    qkv = torch.empty(M, 3 * d, device=dev, dtype=torch.bfloat16)
    xo = torch.empty(M, d, device=dev)
    xa = torch.empty(1, M, d, device=dev, dtype=torch.bfloat16)
    hm = torch.empty(1, M, dm, device=dev, dtype=torch.bfloat16)
    ssq_o = torch.empty(M, d // 64, device=dev)
    att = torch.empty(1, M, d, device=dev, dtype=torch.bfloat16)

    def g_qkv():
        packing.gemm(A512, Wqkv, M, 3 * d, d, P, bias=bq, ssq_in=ssq, rms_dim=d, rms_eps=1e-6, out_bf16=qkv,
                     out_precision=P)

    def g_o():
        packing.gemm(A512, Wo, M, d, d, P, bias=bo, resid=x, out_f32=xo, out_bf16=xa, out_precision=P, ssq_out=ssq_o)

    def g_fc1():
        packing.gemm(A512, W1, M, dm, d, P, bias=b1, ssq_in=ssq, rms_dim=d, rms_eps=1e-6, act=2, out_bf16=hm,
                     out_precision=P)

    def g_fc2():
        packing.gemm(A2048, W2, M, d, dm, P, bias=b2, resid=x, out_f32=xo, out_bf16=xa, out_precision=P,
                     ssq_out=ssq_o)

    n_seq, N = 16, M // 16

    def g_att():
        _lib.check(lib.avd_attention(qkv.data_ptr(), 0, att.data_ptr(), P, n_seq, N, 8, 64, _lib.stream_ptr()))

    Mt = 16 * 37
    P3 = _lib.AVD_PREC_BF16X3
    Ah = mk(Mt, d, 2)
    Wh = mk(d, d, 2)
    bh = torch.randn(d, device=dev)
    yh = torch.empty(Mt, d, device=dev)

    def g_head():
        packing.gemm(Ah, Wh, Mt, d, d, P3, bias=bh, out_f32=yh)

    temb = torch.empty(1, 256, device=dev)
    tt = torch.zeros(1, device=dev)

    def g_tiny():
        _lib.check(lib.avd_timestep_embedding(tt.data_ptr(), temb.data_ptr(), 1, 256, 10000.0, _lib.stream_ptr()))

    yl = torch.randn(8, 512, device=dev)
    wl, bl, ol = torch.ones(512, device=dev), torch.zeros(512, device=dev), torch.empty(8, 512, device=dev)

    def g_tiny_pdl():
        _lib.check(lib.avd_layernorm_act(yl.data_ptr(), wl.data_ptr(), bl.data_ptr(), 8, 512, 1e-5, 0, ol.data_ptr(), None, 0,
                                         _lib.stream_ptr()))

    cases = [("tiny elementwise (launch floor)", g_tiny, 0.0),
             ("tiny elementwise with PDL", g_tiny_pdl, 0.0),
             (f"qkv   {M}x1536x512", g_qkv, 2.0 * M * 1536 * 512),
             (f"attn  16 seq x {N} x 8 heads", g_att, 4.0 * 16 * N * N * 512),
             (f"oproj {M}x512x512", g_o, 2.0 * M * 512 * 512),
             (f"fc1   {M}x2048x512", g_fc1, 2.0 * M * 2048 * 512),
             (f"fc2   {M}x512x2048", g_fc2, 2.0 * M * 512 * 2048),
             (f"head  {Mt}x512x512 bf16x3", g_head, 2.0 * Mt * 512 * 512)]
    if args.once:
        for _ in range(3):
            g_qkv(); g_att(); g_o(); g_fc1(); g_fc2()
        torch.cuda.synchronize()
        return
    for name, fn, flops in cases:
        us = time_graph(fn, args.reps)
        tf = flops / us * 1e-6 if flops else 0.0
        print(f"{name:40s} {us:8.2f} us/launch  {tf:8.1f} TFLOP/s", flush=True)

    def layer():
        g_qkv(); g_att(); g_o(); g_fc1(); g_fc2()
    us = time_graph(layer, 8)
    print(f"{'one block (5 kernels)':40s} {us:8.2f} us", flush=True)

    if args.trace:
        buf = torch.zeros(32 * 512, device=dev, dtype=torch.int64)
        g_att()
        torch.cuda.synchronize()
        _lib.check(lib.avd_debug_trace(buf.data_ptr(), buf.numel()))
        g_att()
        torch.cuda.synchronize()
        _lib.check(lib.avd_debug_trace(None, 0))
        t = buf.view(-1, 16, 2).cpu()
        t = t[t[:, 0, 0] > 0]
        if t.shape[0] == 0:                      # grid too large for the trace buffer: the kernel did not stamp
            t = torch.ones(1, 16, 2, dtype=torch.int64)
        gt = (t[:, :, 0] - t[:, 0, 0].min()).float()
        ck = (t[:, :, 1] - t[:, 0:1, 1]).float()
        an = ["start", "setup done", "iss: Q,K landed", "iss: S issued, K prefetch", "iss: PV0 issued", "sm: S ready",
              "sm: pass1 done", "sm: P written", "sm: O ready", "sm: O accumulated", "sm: stored", "exit"]
        print(f"--- trace attention: {t.shape[0]} CTAs; kernel span {gt[:, 11].max():.0f} ns")
        for i, nm in enumerate(an):
            print(f"   {nm:26s} cycles: median {ck[:, i].median():8.0f} max {ck[:, i].max():8.0f} | ns since kernel start: "
                  f"median {gt[:, i].median():8.0f} max {gt[:, i].max():8.0f}")
        import os
        v4 = _lib.get_option("gemm_v") != 3
        for name, fn in (("qkv", g_qkv), ("oproj", g_o), ("fc1", g_fc1), ("fc2", g_fc2), ("head", g_head)):
            buf = torch.zeros(32 * 512, device=dev, dtype=torch.int64)
            fn()
            torch.cuda.synchronize()
            _lib.check(lib.avd_debug_trace(buf.data_ptr(), buf.numel()))
            fn()
            torch.cuda.synchronize()
            _lib.check(lib.avd_debug_trace(None, 0))
            t = buf.view(-1, 16, 2).cpu()
            live = t[:, 0, 0] > 0
            t = t[live]
            g0 = t[:, 0, 0].min()
            gt = (t[:, :, 0] - g0).float()      # ns since the first CTA started
            ck = (t[:, :, 1] - t[:, 0:1, 1]).float()   # cycles since CTA start
            names = ["start", "setup done", "-", "pipe0: last mma issued", "epi: rinv+resid ready",
                     "epi: acc ready", "epi: stores read", "exit", "ch0: tmem ld done", "ch0: math done",
                     "ch0: sts done", "ch0: fence+syncwarp", "ch0: tma store issued", "-",
                     "setup: barriers initialised", "setup: tmem allocated"]
            if v4:
                names = ["start", "setup done", "iss0: first stage landed", "iss0: last mma issued",
                         "epi: tile0 rinv/resid issued", "epi: tile0 acc ready", "epi: tile0 done", "exit",
                         "epi: last acc ready", "epi: last tile done", "prod0: last load issued"]
            print(f"--- trace {name}: {t.shape[0]} CTAs; kernel span {gt[:, 7].max():.0f} ns "
                  f"(CTA start spread {gt[:, 0].max():.0f} ns)")
            for i, nm in enumerate(names):
                if nm == "-":
                    continue
                print(f"   {nm:20s} cycles since CTA start: median {ck[:, i].median():8.0f}  max {ck[:, i].max():8.0f}"
                      f"   | ns since kernel start: median {gt[:, i].median():8.0f} max {gt[:, i].max():8.0f}")


if __name__ == "__main__":
    main()
