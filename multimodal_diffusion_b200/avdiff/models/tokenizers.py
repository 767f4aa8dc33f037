"""tokenizers.py -- latent <-> token layout classes (mirror of avdiff/models/tokenizers.py).

Thin wrappers over the gather/scatter kernels in ``avdiff.utils.ops``; bit-exact with the reference.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

import torch

from ..utils import ops


@dataclass
class VideoTokenizerCfg:
    tube_t: int = 2
    tube_h: int = 4
    tube_w: int = 4


class VideoTokenizer(torch.nn.Module):
    """Reference: tokenizers.py:33-58.  encode: [B,C,T,H,W] -> [B,N,C*t*h*w]; decode is the inverse."""

    def __init__(self, cfg: VideoTokenizerCfg | None = None):
        super().__init__()
        self.cfg = cfg or VideoTokenizerCfg()

    @property
    def tube(self) -> Tuple[int, int, int]:
        return self.cfg.tube_t, self.cfg.tube_h, self.cfg.tube_w

    def encode(self, z: torch.Tensor) -> torch.Tensor:
        t, h, w = self.tube
        return ops.tube_patch_video(z, t=t, h=h, w=w)

    def decode(self, tokens: torch.Tensor, C: int, T: int, H: int, W: int) -> torch.Tensor:
        t, h, w = self.tube
        return ops.tube_unpatch_video(tokens, C=C, T=T, H=H, W=W, t=t, h=h, w=w)

    def token_dim(self, C_latent: int) -> int:
        t, h, w = self.tube
        return C_latent * t * h * w


@dataclass
class AudioTokenizerCfg:
    chunk_len: int = 4
    stride: int = 4


class AudioTokenizer(torch.nn.Module):
    """Reference: tokenizers.py:68-117.  encode: [B,C,L] -> [B,N,C*chunk_len]; decode: overlap-add back to
    [B,C,L_total] (cropped / zero-padded), one kernel instead of the reference's B*C*N slice updates."""

    def __init__(self, cfg: AudioTokenizerCfg | None = None):
        super().__init__()
        self.cfg = cfg or AudioTokenizerCfg()

    @property
    def chunk(self) -> Tuple[int, int]:
        return self.cfg.chunk_len, self.cfg.stride

    def encode(self, z: torch.Tensor) -> torch.Tensor:
        length, stride = self.chunk
        return ops.latents_to_tokens_audio(z, length, stride)

    def decode(self, tokens: torch.Tensor, C: int, L_total: int) -> torch.Tensor:
        length, stride = self.chunk
        D = tokens.shape[-1]
        assert D == C * length, f"token dim {D} != C*chunk_len {C*length}"
        return ops.tokens_to_latents_audio(tokens, C, length, L_total, stride)

    def token_dim(self, C_latent: int) -> int:
        return C_latent * self.cfg.chunk_len
