"""Chained GEMM launch (avd_gemm_chain.cu) against the one-kernel-per-GEMM schedule on one DDIM+CFG step and on a
50-step graph (development tool): max / rel-L2 differences, run-to-run determinism, flag block left zero."""
from __future__ import annotations

import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from multimodal_diffusion_b200 import _lib  # noqa: E402
from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig  # noqa: E402
from _common import randn, rel_l2, seeded_modules  # noqa: E402


def main() -> None:
    dev = torch.device("cuda:0")
    mods = seeded_modules()
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    target = sys.argv[2] if len(sys.argv) > 2 else "audio"
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 50
    out = {}
    for chain in (0, 1, 1):
        _lib.set_option("gemm_chain", chain)
        fs = FusedSampler(HotPathConfig(sampler_steps=steps, guidance=3.5), mods[0], mods[1], mods[2], mods[3], target, B,
                          device=dev)
        zp = randn(1, *fs.prompt_shape).to(dev)
        z0 = randn(2, *fs.latent_shape).to(dev)
        fs.prepare(zp)
        z = z0.clone()
        tap = torch.empty(2 * B, fs.Nt, fs.Dt, device=dev)
        fs.step(0, z, tap)
        torch.cuda.synchronize()
        fs.check_errors()
        zf = fs.sample_eager(zp, z0)
        torch.cuda.synchronize()
        fs.check_errors()
        zg = fs.sample(zp, z0)
        torch.cuda.synchronize()
        fs.check_errors()
        key = f"chain={chain}" + ("" if f"chain={chain}" not in out else " (again)")
        out[key] = (tap.clone(), z.clone(), zf.clone(), zg.clone())
        print(key, "eps finite", bool(torch.isfinite(tap).all()), "graph == eager", bool(torch.equal(zf, zg)), flush=True)
    a, b, c = out["chain=0"], out["chain=1"], out["chain=1 (again)"]
    print("step-0 eps  chain vs unchained rel-L2", rel_l2(b[0].cpu(), a[0].cpu()))
    print("step-0 z    chain vs unchained rel-L2", rel_l2(b[1].cpu(), a[1].cpu()))
    print(f"{steps}-step z  chain vs unchained rel-L2", rel_l2(b[2].cpu(), a[2].cpu()))
    print("chain run-to-run bit-identical:", bool(torch.equal(b[2], c[2])), bool(torch.equal(b[0], c[0])))


if __name__ == "__main__":
    main()
