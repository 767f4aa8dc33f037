"""multimodal_diffusion_b200 -- B200 (sm_100a) implementation of avdiff's DDIM + CFG sampling hot path.

Product code only: the CUDA kernels + C ABI (csrc/, include/avdiff_b200.h), the ctypes binding (_lib)
and the host-side mirror of the reference's Python surface (avdiff/).  Nothing here imports ``oracle/``.
"""
__version__ = "0.1.0"
