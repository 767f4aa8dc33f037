"""Per-kernel counts of the Blackwell-specific SASS instructions in libavdiff_b200.so (development tool).

    python tools/sass_summary.py [--out profiles/r2_sass_summary.txt]

Runs `cuobjdump -sass` on the built library (works in the CPU-only build container) and, for every kernel, counts
the mnemonics that prove the tcgen05 / TMEM / TMA path: UTCHMMA (tcgen05.mma, `.2CTA` = cta_group::2), LDTM / STTM
(tcgen05.ld / st), UTMALDG / UTMASTG (TMA tensor load / store, `.2CTA` multicast-to-pair form), UTCBAR
(tcgen05.commit -> mbarrier), UTCCP (tcgen05.cp), SYNCS (mbarrier ops), UBLKCP (cp.async.bulk), plus registers from
`cuobjdump -res-usage`.  Kernels without any of them (elementwise / codec kernels) are listed by name only.
"""
from __future__ import annotations

import argparse
import re
import subprocess
import sys
from collections import Counter, OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "multimodal_diffusion_b200" / "libavdiff_b200.so"
KEYS = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMALDG.2CTA", "UTMASTG", "UTCBAR", "UTCCP", "SYNCS", "UBLKCP"]


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout
        return out.splitlines()
    except Exception:
        return list(names)


def short(name: str) -> str:
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    return name.replace("avd::", "")


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default=str(LIB))
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    sass = subprocess.run(["cuobjdump", "-sass", args.lib], capture_output=True, text=True, check=True).stdout
    per: "OrderedDict[str, Counter]" = OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per[cur] = Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if not m:
            continue
        op = m.group(1)
        per[cur]["_n"] += 1
        base = op.split(".")[0]
        if base == "UTCHMMA":
            per[cur]["UTCHMMA.2CTA" if ".2CTA" in op else "UTCHMMA"] += 1
        elif base == "UTMALDG":
            per[cur]["UTMALDG.2CTA" if ".2CTA" in op else "UTMALDG"] += 1
        elif base in ("LDTM", "STTM", "UTMASTG", "UTCBAR", "UTCCP", "SYNCS", "UBLKCP"):
            per[cur][base] += 1
    res = subprocess.run(["cuobjdump", "-res-usage", args.lib], capture_output=True, text=True).stdout
    regs = {}
    fn = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            fn = m.group(1)
        m = re.search(r"REG:(\d+)", line)
        if m and fn:
            regs[fn] = int(m.group(1))
    names = list(per.keys())
    dem = [short(d) for d in demangle(names)]
    lines = [f"SASS summary of {Path(args.lib).name} (cuobjdump -sass, sm_100a); counts are static instructions per kernel",
             f"{'kernel':78s} {'regs':>4s} {'instr':>6s} " + " ".join(f"{k:>12s}" for k in KEYS)]
    tot = Counter()
    plain = []
    for n, d in sorted(zip(names, dem), key=lambda t: t[1]):
        c = per[n]
        if not any(c[k] for k in KEYS):
            plain.append(d)
            continue
        for k in KEYS:
            tot[k] += c[k]
        lines.append(f"{d[:78]:78s} {regs.get(n, 0):4d} {c['_n']:6d} " + " ".join(f"{c[k]:12d}" for k in KEYS))
    lines.append(f"{'TOTAL (' + str(len(names)) + ' kernels)':78s} {'':4s} {'':6s} " + " ".join(f"{tot[k]:12d}" for k in KEYS))
    lines.append("")
    lines.append("kernels without tensor-core / TMA / mbarrier instructions (CUDA-core byte movers, codecs, norms):")
    for d in sorted(set(plain)):
        lines.append("  " + d[:150])
    text = "\n".join(lines) + "\n"
    if args.out:
        Path(args.out).write_text(text)
        print(f"wrote {args.out}: {len(names)} kernels, totals " + ", ".join(f"{k} {tot[k]}" for k in KEYS if tot[k]))
    else:
        sys.stdout.write(text)


if __name__ == "__main__":
    main()
