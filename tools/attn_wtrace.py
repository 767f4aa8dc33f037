"""Per-warp timeline of one CTA of the multi-block attention kernel (development tool).

    tools/exp_build.sh wtrace -DAVD_DEBUG_HOOKS -DAVD_ATT_WTRACE
    AVD_LIB_PATH=build/exp/wtrace/libavdiff_b200.so python tools/attn_wtrace.py [--N 2048] [--seq 16]

Needs a library built with -DAVD_ATT_WTRACE: lane 0 of each of the 16 softmax warps of CTA 0 stamps clock64 at 9 points
of key blocks 3, 4, 5 (csrc/avd_attn.cu ATT_WMARK).  Prints, per warp, the cycles between the stamps of block 4 and the
block period; the grid is the full one, so MUFU / TMEM / L2 contention is that of the timed kernel.
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from multimodal_diffusion_b200 import _lib  # noqa: E402

NAMES = ["top", "S ready", "pass done", "max exchanged", "repeat done", "P arrived", "MMAs issued", "O ready", "O read"]


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--N", type=int, default=2048)
    ap.add_argument("--seq", type=int, default=16)
    ap.add_argument("--bf16", action="store_true")
    ap.add_argument("--blk", type=int, default=1, help="0..2: which of the key blocks 3, 4, 5 to print")
    args = ap.parse_args()
    lib = _lib.load()
    dev = torch.device("cuda:0")
    H, hd = 8, 64
    dt = torch.bfloat16 if args.bf16 else torch.float16
    g = torch.Generator().manual_seed(1)
    qkv = (torch.randn(args.seq * args.N, 3 * H * hd, generator=g) * 1.5).to(dt).to(dev)
    out = torch.zeros(1, args.seq * args.N, H * hd, device=dev, dtype=dt)
    prec = _lib.AVD_PREC_BF16 if args.bf16 else _lib.AVD_PREC_F16

    def run():
        _lib.check(lib.avd_attention(qkv.data_ptr(), 0, out.data_ptr(), prec, args.seq, args.N, H, hd, _lib.stream_ptr()))
        torch.cuda.synchronize()

    run()
    n_ctas = ((args.N + 255) // 256) * H * args.seq
    buf = torch.zeros(max(32 * n_ctas, 1024), device=dev, dtype=torch.int64)
    _lib.check(lib.avd_debug_trace(buf.data_ptr(), buf.numel()))
    run()
    _lib.check(lib.avd_debug_trace(None, 0))
    t = buf[:17 * 3 * 12].view(17, 3, 12).cpu()
    if int(t.max()) == 0:
        print("no stamps: library not built with -DAVD_ATT_WTRACE, or the single-block kernel ran")
        return
    B = args.blk
    t0 = int(t[:, B, 0][t[:, B, 0] > 0].min())
    print(f"attention {args.seq} seq x {args.N} keys x {H} heads, CTA 0, key block 4; cycles since the first warp's block top")
    print("warp tile half quad | " + " ".join(f"{n:>13s}" for n in NAMES) + " |  period(3->4) period(4->5)")
    if _lib.get_option("attn_fa") != 0:
        names = ["top", "S ready", "pass done", "max exchanged", "rescale/repeat", "P arrived"]
        print("warp tile half quad | " + " ".join(f"{n:>14s}" for n in names) + " |  period(3->4) period(4->5)")
        for w in range(8 if _lib.get_option("attn_fa") in (2, 3) else 16):
            if int(t[w, B, 0]) == 0:
                continue
            row = " ".join(f"{int(t[w, B, sl]) - t0:14d}" for sl in range(6))
            print(f"{w:4d} {w >> 3:4d} {(w >> 2) & 1:4d} {w & 3:4d} | {row} | {int(t[w, B, 0] - t[w, 0, 0]):13d} "
                  f"{int(t[w, 2, 0] - t[w, B, 0]):12d}")
        print("inside the pass:    chunk 0: ld done, exps done, P stored | chunk 1: ld done, exps done, P stored")
        for w in range(0 if _lib.get_option("attn_fa") >= 2 else 16):
            if int(t[w, B, 0]) == 0:
                continue
            print(f"{w:4d}                | " + " ".join(f"{int(t[w, B, sl]) - t0:10d}" for sl in range(6, 12)))
        cw = 8 if _lib.get_option("attn_fa") in (2, 3) else 16
        c = [int(t[cw, B, sl]) - t0 for sl in range(6)]
        print(f"control warp: tile 0: S read seen {c[0]}, S(j+1) issued {c[1]}, P V(j) issued {c[2]} | tile 1: S read seen {c[3]}, "
              f"S(j+1) issued {c[4]}, P V(j) issued {c[5]} | period {int(t[cw, 2, 0] - t[cw, 1, 0])}")
        return
    for w in range(16):
        if int(t[w, B, 0]) == 0:
            continue
        row = []
        for sl in range(9):
            v = int(t[w, B, sl])
            row.append(f"{v - t0:13d}" if v > 0 else f"{'-':>13s}")
        p34 = int(t[w, B, 0] - t[w, 0, 0])
        p45 = int(t[w, 2, 0] - t[w, B, 0])
        print(f"{w:4d} {w >> 3:4d} {(w >> 2) & 1:4d} {w & 3:4d} | " + " ".join(row) + f" | {p34:13d} {p45:12d}")
        if (w & 7) == 0:                      # the tile's MMA issuer
            a, b, c, d = (int(t[w, B, sl]) - t0 for sl in (9, 10, 11, 6))
            print(f"      issuer: P/V/R waits over {a}, PV issued+committed {b}, K(j+1) there {c}, S(j+1) issued+committed {d}")


if __name__ == "__main__":
    main()
