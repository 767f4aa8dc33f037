"""Stress the persistent GEMM inside the fused sampler (development tool): replays the V2A B=8 sampling graph and an
A2V B=64 call, then prints the barrier-timeout record."""
import ctypes as C
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from multimodal_diffusion_b200 import _lib  # noqa: E402
from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig  # noqa: E402
from _common import randn, seeded_modules  # noqa: E402

dev = torch.device("cuda:0")
lib = _lib.load()
mods = seeded_modules()


def report(where):
    try:
        torch.cuda.synchronize()
    except Exception as ex:
        out = (C.c_uint * 256)()
        lib.avd_debug_gemm_timeout(out)
        print(where, "CUDA FAILURE", str(ex).splitlines()[0], "timeout records:", out[0], [list(out[8 + 8 * i: 16 + 8 * i]) for i in range(min(out[0], 30))], flush=True)
        raise SystemExit(1)
    out = (C.c_uint * 256)()
    lib.avd_debug_gemm_timeout(out)
    print(where, "timeout records:", out[0], [list(out[8 + 8 * i: 16 + 8 * i]) for i in range(min(out[0], 30))], flush=True)


for target, B, reps in (("audio", 8, int(sys.argv[1]) if len(sys.argv) > 1 else 30), ("video", 64, 3)):
    fs = FusedSampler(HotPathConfig(sampler_steps=50, guidance=3.5), mods[0], mods[1], mods[2], mods[3], target, B, device=dev)
    zp = randn(1, *fs.prompt_shape).to(dev)
    z = randn(2, *fs.latent_shape).to(dev)
    for r in range(reps):
        try:
            out = fs.sample(zp, z)
        except Exception:
            report(f"{target} B={B} rep {r} (sample raised)")
        if r % 10 == 0:
            report(f"{target} B={B} rep {r}")
    report(f"{target} B={B} before final check")
    report(f"{target} B={B} done; finite={bool(torch.isfinite(out).all())}")
