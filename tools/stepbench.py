"""Whole-step timing on one B200 (development tool): one DDIM+CFG step of the fused sampler in a CUDA graph, and its
parts (core blocks only; everything else), V2A B = 8 at mvp shapes."""
from __future__ import annotations

import ctypes as C
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from multimodal_diffusion_b200 import _lib  # noqa: E402
from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig  # noqa: E402
from _common import randn, seeded_modules  # noqa: E402
from kbench import time_graph  # noqa: E402


def main() -> None:
    dev = torch.device("cuda:0")
    lib = _lib.load()
    mods = seeded_modules()
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    target = sys.argv[2] if len(sys.argv) > 2 else "audio"
    for a in sys.argv:                      # --opt name=value sets a library tuning switch (e.g. --opt chains=2)
        if a.startswith("--opt="):
            k, v = a[6:].split("=")
            _lib.set_option(k, int(v))
    kw = dict(Fa=500, video_latent=(40, 16, 16)) if "--cfg10" in sys.argv else {}   # the 10 s clip: N = 445 joint tokens
    fs = FusedSampler(HotPathConfig(sampler_steps=50, guidance=3.5, **kw), mods[0], mods[1], mods[2], mods[3], target, B,
                      device=dev)
    zp = randn(1, *fs.prompt_shape).to(dev)
    z = randn(2, *fs.latent_shape).to(dev)
    fs.prepare(zp)
    torch.cuda.synchronize()

    def step():
        _lib.check(lib.avd_sample_step(C.byref(fs.args), 3, z.data_ptr(), None, _lib.stream_ptr()))

    us = time_graph(step, 10)
    opts = " ".join(a[6:] for a in sys.argv if a.startswith("--opt="))
    print(f"one DDIM+CFG step ({target}, B={B}) {opts}: {us:9.2f} us  -> {B / (us * 50e-6):8.1f} clips/s")
    if "--kinds" not in sys.argv:
        return
    # per-kind breakdown with the event profiler (eager; includes launch gaps, use for shares only)
    lib.avd_prof_reset(); lib.avd_prof_enable(1)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    lib.avd_prof_enable(0)
    for k in range(lib.avd_prof_num_kinds()):
        ms, n = C.c_double(), C.c_int64()
        lib.avd_prof_collect(k, C.byref(ms), C.byref(n))
        if n.value:
            print(f"   {lib.avd_prof_kind_name(k).decode():16s} {n.value // 5:4d} launches/step  {ms.value * 1e3 / n.value:8.2f} us each (eager)")


if __name__ == "__main__":
    main()
