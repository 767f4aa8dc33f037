"""One step with the chained launch; on failure print the barrier-timeout records (development tool)."""
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
try:
    fs.prepare(zp)
    fs.step(0, z, None)
    torch.cuda.synchronize()
    print("step ok", bool(torch.isfinite(z).all()))
except Exception as ex:
    print("FAILED:", str(ex).splitlines()[0])
rec = (C.c_uint * 256)()
lib.avd_debug_gemm_timeout(rec)
print("timeouts:", rec[0])
for i in range(min(rec[0], 30)):
    r = [rec[8 + i * 8 + k] for k in range(8)]
    print("  tag", r[0], "cta", r[1], "thread", r[2], "phase", r[3], "parity", r[4], r[5:])
