"""schedules.py -- per-modality diffusion schedule object (mirror of avdiff/models/schedules.py)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Literal

import torch

from ..utils import schedule_utils as su

ScheduleKind = Literal["linear", "cosine", "sigmoid"]


@dataclass
class ModalitySchedule:
    """Reference: schedules.py:26-87.  betas / alphas / alpha-bar tables plus forward and reverse ops."""
    kind: ScheduleKind
    steps: int
    betas: torch.Tensor
    alphas: torch.Tensor
    alphas_cumprod: torch.Tensor

    @classmethod
    def make(cls, *, kind: ScheduleKind = "cosine", steps: int = 1000, min_beta: float = 1e-4,
             max_beta: float = 2e-2) -> "ModalitySchedule":
        betas = su.make_beta_schedule(steps=steps, kind=kind, min_beta=min_beta, max_beta=max_beta)
        alphas, alpha_bar = su.alphas_cumprod_from_betas(betas)
        return cls(kind=kind, steps=int(steps), betas=betas, alphas=alphas, alphas_cumprod=alpha_bar)

    def to(self, device: torch.device) -> "ModalitySchedule":
        self.betas, self.alphas = self.betas.to(device), self.alphas.to(device)
        self.alphas_cumprod = self.alphas_cumprod.to(device)
        return self

    def q_sample(self, z0: torch.Tensor, t: torch.Tensor, noise: torch.Tensor | None = None):
        """The reference passes ``noise=`` to a function whose keyword is ``eps=`` and raises TypeError
        (schedules.py:61); here the call goes through."""
        return su.q_sample(z0, t, self.alphas_cumprod, eps=noise)

    def ddim_step(self, z_t: torch.Tensor, t: torch.Tensor, t_prev: torch.Tensor, eps_hat: torch.Tensor,
                  eta: float = 0.0) -> torch.Tensor:
        """Reference: schedules.py:65-74 -> avd_ddim_step."""
        return su.ddim_step(z_t, t, t_prev, eps_hat, self.alphas_cumprod, eta=eta)

    def make_sampling_schedule(self, steps_sample: int) -> torch.Tensor:
        return su.make_sampling_schedule(self.steps, steps_sample)

    def timestep_embedding(self, t: torch.Tensor, dim: int, max_period: int = 10_000) -> torch.Tensor:
        return su.timestep_embedding(t, dim=dim, max_period=max_period)


def build_schedules_from_config(cfg: Dict) -> Dict[str, ModalitySchedule]:
    """Reference: schedules.py:90-109."""
    out = {}
    for m in ("video", "audio"):
        c = cfg["diffusion"][m]
        out[m] = ModalitySchedule.make(kind=c.get("schedule", "cosine"), steps=int(c.get("steps", 1000)),
                                       min_beta=float(c.get("min_beta", 1e-4)), max_beta=float(c.get("max_beta", 2e-2)))
    return out
