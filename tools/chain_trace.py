"""Phase timeline of one chained launch (debug-hooks build): medians over CTAs, cycles since the CTA's producer entered phase 0."""
import ctypes as C, sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from multimodal_diffusion_b200 import _lib
from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig
from _common import randn, seeded_modules
dev = torch.device("cuda:0")
mods = seeded_modules()
lib = _lib.load()
fs = FusedSampler(HotPathConfig(sampler_steps=50, guidance=3.5), mods[0], mods[1], mods[2], mods[3], "audio", 8, device=dev)
zp = randn(1, *fs.prompt_shape).to(dev); z = randn(2, *fs.latent_shape).to(dev)
fs.prepare(zp)
fs.step(0, z, None); torch.cuda.synchronize()
buf = torch.zeros(148 * 4 * 16, device=dev, dtype=torch.int64)
_lib.check(lib.avd_debug_trace(buf.data_ptr(), buf.numel()))
fs.step(1, z, None); torch.cuda.synchronize()
_lib.check(lib.avd_debug_trace(None, 0))
t = buf.view(148, 4, 16).cpu()[:136].double()      # the FIRST chained launch of the step (block 0)
t0 = t[:, 0:1, 0:1]            # clock64 is per SM: every CTA's stamps relative to its own first one
names = ["producer enters", "weights issued", "dependency seen", "last load issued", "acc complete", "stores issued", "signalled",
         "ch0 tmem ld done", "ch0 resid added", "ch0 staged", "ch0 pair barrier", "ch0 f32 stored", "ch0 16-bit stored"]
order = [0, 1, 2, 3, 4, 7, 8, 9, 10, 11, 12, 5, 6]
for ph in range(4):
    if t[:, ph, 0].max() == 0:
        continue
    print(f"phase {ph}")
    for i in order:
        nm = names[i]
        v = t[:, ph, i]
        if v.max() == 0:
            continue
        v = v - t0[:, 0, 0]
        print(f"   {nm:18s} median {v.median():8.0f} cycles   min {v.min():8.0f}  max {v.max():8.0f}")
