"""schedule_utils.py -- noise schedules, timestep embedding and the DDIM update (mirror of
avdiff/utils/schedule_utils.py).

Host side (once per sampling call, a few thousand scalars): beta / alpha-bar tables and the integer
sampling schedule, evaluated with the same fp32 torch-CPU operations the reference uses so the tables are
bit-identical.  Device side: ``timestep_embedding`` and ``ddim_step`` run as kernels of libavdiff_b200.so.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

from ... import _lib


# ------------------------------------------------------------------ host tables

def make_beta_schedule(steps: int, kind: str = "cosine", min_beta: float = 1e-4, max_beta: float = 2e-2) -> torch.Tensor:
    """Reference: schedule_utils.py:14-49.  betas[t], t = 0..steps-1, fp32, clamped to [1e-8, 0.999]."""
    name = kind.lower()
    if name == "cosine":
        s = 0.008
        grid = torch.linspace(0, steps, steps + 1, dtype=torch.float32)
        f = torch.cos(((grid / steps + s) / (1 + s)) * math.pi / 2) ** 2
        a_bar = f / f[0]
        return (1 - (a_bar[1:] / a_bar[:-1])).clamp(1e-8, 0.999)
    if name == "linear":
        return torch.linspace(min_beta, max_beta, steps, dtype=torch.float32).clamp(1e-8, 0.999)
    if name == "sigmoid":
        ramp = torch.sigmoid(torch.linspace(-6, 6, steps, dtype=torch.float32))
        return (min_beta + (max_beta - min_beta) * ramp).clamp(1e-8, 0.999)
    raise ValueError(f"Unknown schedule kind: {kind}")


def alphas_cumprod_from_betas(betas: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Reference: schedule_utils.py:52-57."""
    alphas = 1.0 - betas.to(dtype=torch.float32)
    return alphas, torch.cumprod(alphas, dim=0)


def make_sampling_schedule(T_train: int, T_sample: int) -> torch.Tensor:
    """Reference: schedule_utils.py:132-143.  round(linspace(T_train-1, -1, T_sample+1)) as int64, capped at its
    first entry."""
    sched = torch.round(torch.linspace(T_train - 1, -1, T_sample + 1)).to(torch.long)
    return torch.minimum(sched, torch.full_like(sched, int(sched[0])))


def timestep_embedding_host(timesteps: torch.Tensor, dim: int, max_period: int = 10000) -> torch.Tensor:
    """The reference formula (schedule_utils.py:64-86) on the host, used once per sampling call to build the
    [steps, dim] table the fused input builder reads."""
    t = timesteps.detach().cpu()
    if t.dtype not in (torch.float32, torch.float64):
        t = t.float()
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32) / half)
    ang = t.unsqueeze(1) * freqs.unsqueeze(0)
    emb = torch.cat([torch.cos(ang), torch.sin(ang)], dim=1)
    if dim % 2 == 1:
        emb = torch.nn.functional.pad(emb, (0, 1))
    return emb


# ------------------------------------------------------------------ device ops

def timestep_embedding(timesteps: torch.Tensor, dim: int, max_period: int = 10000) -> torch.Tensor:
    """Reference: schedule_utils.py:64-86.  [B] -> [B, dim] = [cos(t f_k) | sin(t f_k)].  C ABI: avd_timestep_embedding."""
    if not timesteps.is_cuda:
        raise _lib.AvdError("timesteps must live on a CUDA device (no CPU path)")
    t = timesteps if timesteps.dtype == torch.float32 else timesteps.float()
    t = t.contiguous()
    B = t.shape[0]
    emb = torch.empty(B, dim, device=t.device, dtype=torch.float32)
    if emb.numel() > 0:
        _lib.check(_lib.load().avd_timestep_embedding(t.data_ptr(), emb.data_ptr(), B, dim, float(max_period),
                                                      _lib.stream_ptr()))
    return emb


def _gather(a_bar: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
    return a_bar[t.long()]


def q_sample(x0: torch.Tensor, t: torch.Tensor, alpha_bar: torch.Tensor, eps: Optional[torch.Tensor] = None):
    """Reference: schedule_utils.py:104-125 (training-time forward noising; NOT on the sampling hot path, kept
    for API completeness with plain tensor ops)."""
    a = _gather(alpha_bar.to(x0.device), t.to(x0.device))
    a = a.reshape([-1] + [1] * (x0.dim() - 1))
    if eps is None:
        eps = torch.randn_like(x0)
    return torch.sqrt(a) * x0 + torch.sqrt(torch.clamp(1.0 - a, min=0.0)) * eps, eps


def ddim_step(x_t: torch.Tensor, t_now: torch.Tensor, t_prev: torch.Tensor, eps_hat: torch.Tensor,
              alpha_bar: torch.Tensor, eta: float = 0.0) -> torch.Tensor:
    """Reference: schedule_utils.py:146-200.  One DDIM update; alpha_bar[-1] := 1.  C ABI: avd_ddim_step.
    The arithmetic follows the reference's fp32 operation order exactly (bit-exact for eta = 0)."""
    x = _lib.require_cuda_f32(x_t, "x_t")
    e = _lib.require_cuda_f32(eps_hat, "eps_hat")
    if e.shape != x.shape:
        raise RuntimeError(f"eps_hat shape {tuple(e.shape)} does not match x_t {tuple(x.shape)}")
    B = x.shape[0]
    dev = x.device
    tn = t_now.to(device=dev, dtype=torch.long).contiguous()
    tp = t_prev.to(device=dev, dtype=torch.long).contiguous()
    if tn.numel() != B or tp.numel() != B:
        raise RuntimeError("t_now / t_prev must have one entry per batch element")
    ab = alpha_bar.to(device=dev, dtype=torch.float32).contiguous()
    if B and (int(tn.max()) >= ab.numel() or int(tp.max()) >= ab.numel()):
        raise IndexError("timestep out of range of alpha_bar")
    noise = torch.randn_like(x) if eta > 0.0 else None
    out = torch.empty_like(x)
    if x.numel() > 0:
        _lib.check(_lib.load().avd_ddim_step(x.data_ptr(), e.data_ptr(), _lib.ptr(noise), out.data_ptr(),
                                             tn.data_ptr(), tp.data_ptr(), ab.data_ptr(), ab.numel(), B,
                                             x.numel() // B, float(eta), _lib.stream_ptr()))
    return out
