"""adapters.py -- token adapters and embeddings (mirror of avdiff/models/adapters.py).

``LinearAdapter`` is the only piece the sampling hot path executes (P_v / P_a, sample_clip.py:48-56,91-92); its
forward runs the tcgen05 GEMM of libavdiff_b200.so in fp32-split mode (inside the fused sampler the adapter is
part of the input-builder kernel instead).  The modality / positional / timestep embedders are API surface the
reference never imports on this path: they are small additive tables kept as plain
tensor code so callers that construct them keep working.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Literal, Optional, Tuple

import numpy as np

import torch
import torch.nn as nn

from ... import _lib, packing
from ..utils import schedule_utils as su
from .mmdt import _no_grad_only, default_precision


class LinearAdapter(nn.Module):
    """Reference: adapters.py:28-38 (xavier init) == sample_clip.py:48-56 (default init; pass xavier=False).
    forward(x [B,N,d_in]) -> [B,N,d_out]."""

    def __init__(self, d_in: int, d_out: int, xavier: bool = True):
        super().__init__()
        self.proj = nn.Linear(d_in, d_out)
        if xavier:
            nn.init.xavier_uniform_(self.proj.weight)
            nn.init.zeros_(self.proj.bias)
        self.precision = default_precision()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _no_grad_only(self)
        xc = _lib.require_cuda_f32(x, "x")
        d_in, d_out = self.proj.in_features, self.proj.out_features
        if xc.shape[-1] != d_in:
            raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied (last dim {xc.shape[-1]} != {d_in})")
        M = xc.numel() // d_in
        prec = _lib.precision_code(self.precision)
        Kp, Np = (d_in + 63) // 64 * 64, (d_out + 31) // 32 * 32
        W = self.proj.weight.detach()
        b = self.proj.bias.detach()
        x2 = xc.reshape(M, d_in)
        if Kp != d_in or Np != d_out:      # zero-pad to the GEMM's tile granularity (exact: adds 0-products)
            Wp = W.new_zeros(Np, Kp); Wp[:d_out, :d_in] = W; W = Wp
            bp = b.new_zeros(Np); bp[:d_out] = b; b = bp
            xp = x2.new_zeros(M, Kp); xp[:, :d_in] = x2; x2 = xp
        y = torch.empty(M, Np, device=xc.device, dtype=torch.float32)
        if M > 0:
            xa, _ = packing.rows_prepare(x2, prec)
            packing.gemm(xa, packing.pack_linear_weight(W, prec), M, Np, Kp, prec, bias=b, out_f32=y)
        return y[:, :d_out].reshape(*xc.shape[:-1], d_out)


class ModalityEmbedding(nn.Module):
    """Reference: adapters.py:41-52.  x + learned per-modality vector."""

    def __init__(self, d: int, modalities=("video", "audio")):
        super().__init__()
        self.lookup = nn.Embedding(len(modalities), d)
        nn.init.normal_(self.lookup.weight, mean=0.0, std=0.02)
        self.index = {m: i for i, m in enumerate(modalities)}

    def forward(self, x: torch.Tensor, modality: str) -> torch.Tensor:
        return x + self.lookup.weight[self.index[modality]].to(x.device, x.dtype).view(1, 1, -1)


def _sin_table(n: int, d: int, device, dtype) -> torch.Tensor:
    pos = torch.arange(n, device=device, dtype=dtype).unsqueeze(1)
    div = torch.exp(torch.arange(0, d, 2, device=device, dtype=dtype) * (-math.log(10000.0) / d))
    table = torch.zeros(n, d, device=device, dtype=dtype)
    table[:, 0::2] = torch.sin(pos * div)
    table[:, 1::2] = torch.cos(pos * div)
    return table


class PositionalEmbedding1D(nn.Module):
    """Reference: adapters.py:65-83.  forward(B, N, device, dtype) -> [B,N,d]."""

    def __init__(self, d: int, max_len: int = 4096, mode: Literal["learned", "sin"] = "learned"):
        super().__init__()
        self.d, self.mode, self.max_len = d, mode, max_len
        if mode == "learned":
            self.table = nn.Embedding(max_len, d)
            nn.init.normal_(self.table.weight, mean=0.0, std=0.02)

    def forward(self, B: int, N: int, device, dtype) -> torch.Tensor:
        pe = self.table(torch.arange(N, device=device)) if self.mode == "learned" \
            else _sin_table(N, self.d, device, dtype)
        return pe.unsqueeze(0).expand(B, N, -1)


class PositionalEmbedding3D(nn.Module):
    """Reference: adapters.py:86-127.  forward(B, Tt, Hh, Ww, device, dtype) -> [B, Tt*Hh*Ww, d], raster order."""

    def __init__(self, d: int, max_T: int = 256, max_H: int = 256, max_W: int = 256,
                 mode: Literal["learned", "sin"] = "learned"):
        super().__init__()
        self.d, self.mode = d, mode
        if mode == "learned":
            self.t_table, self.h_table, self.w_table = nn.Embedding(max_T, d), nn.Embedding(max_H, d), nn.Embedding(max_W, d)
            for tab in (self.t_table, self.h_table, self.w_table):
                nn.init.normal_(tab.weight, mean=0.0, std=0.02)

    def forward(self, B: int, Tt: int, Hh: int, Ww: int, device, dtype) -> torch.Tensor:
        N = Tt * Hh * Ww
        if self.mode == "learned":
            pe = (self.t_table(torch.arange(Tt, device=device))[:, None, None, :]
                  + self.h_table(torch.arange(Hh, device=device))[None, :, None, :]
                  + self.w_table(torch.arange(Ww, device=device))[None, None, :, :]).view(N, self.d)
        else:
            pe = _sin_table(N, self.d, device, dtype)
        return pe.unsqueeze(0).expand(B, -1, -1)


@dataclass
class TimestepCfg:
    dim: int = 256
    mode: Literal["sin", "mlp"] = "sin"


class TimestepEmbedder(nn.Module):
    """Reference: adapters.py:137-158.  t[B] -> [B, dim]; the sinusoidal base runs as a kernel for CUDA inputs."""

    def __init__(self, cfg: TimestepCfg):
        super().__init__()
        self.cfg = cfg
        self.mlp = nn.Sequential(nn.Linear(cfg.dim, cfg.dim * 2), nn.SiLU(), nn.Linear(cfg.dim * 2, cfg.dim)) \
            if cfg.mode == "mlp" else None

    def forward(self, t: torch.Tensor) -> torch.Tensor:
        base = su.timestep_embedding(t, dim=self.cfg.dim) if t.is_cuda else su.timestep_embedding_host(t, self.cfg.dim)
        return self.mlp(base) if self.mlp is not None else base


@torch.no_grad()
def token_additive_table(*, modality_embedding: Optional["ModalityEmbedding"] = None, modality: Optional[str] = None,
                         positional=None, grid: Tuple[int, ...] = ()) -> torch.Tensor:
    """The per-token rows [N, d] that `ModalityEmbedding` (adapters.py:41-52) and a positional embedding
    (adapters.py:65-127) ADD to the adapted tokens of one modality, summed once on the host: the table
    ``FusedSampler(add_v=..., add_a=...)`` hands to the fused input builder, which adds it in the adapter's epilogue
     instead of running two more elementwise passes per step.
    grid: (Tt, Hh, Ww) for PositionalEmbedding3D, (N,) for PositionalEmbedding1D."""
    table = None
    if positional is not None:
        dev = next((p.device for p in positional.parameters()), torch.device("cpu"))
        pe = positional(1, *grid, dev, torch.float32)[0]
        table = pe.detach().clone().float()
    if modality_embedding is not None:
        vec = modality_embedding.lookup.weight[modality_embedding.index[modality]].detach().float()
        n = int(table.shape[0]) if table is not None else int(np.prod(grid))
        # same association as the modules applied in sequence: (x + modality) + positional
        table = vec.view(1, -1).expand(n, -1).clone() if table is None else vec.view(1, -1).to(table.device) + table
    if table is None:
        raise ValueError("give a modality embedding and / or a positional embedding")
    return table.contiguous()
