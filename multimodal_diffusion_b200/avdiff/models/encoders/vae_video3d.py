"""vae_video3d.py -- mirror of avdiff/models/encoders/vae_video3d.py: video <-> 3-D latent grid.

Same constructor order, parameter names (``enc_net.{i}.{0,2}``, ``to_lat`` | ``to_mu`` / ``to_logv``, ``from_lat``,
``dec_net.{i}.{0,2}``, ``to_img``) and default init as the reference, so its ``state_dict`` loads strictly; ``encode`` /
``decode`` run in libavdiff_b200.so (csrc/avd_vae.cu), fp32, GroupNorm statistics reduced deterministically.

    x [B,3,T,H,W] --encode--> z [B,Cv,T//t_down,H//s_down,W//s_down]     (non-divisible inputs are center-cropped)
    z --decode--> x_hat [B,3,T,H,W] in [0,1] (sigmoid) or [-1,1] (tanh)

The reference's sampler hands ``encode`` a [1,T,3,H,W] tensor (sample_clip.py:306), which its own encoder
rejects; the mirror's ``sample_one_direction`` passes the documented [1,3,T,H,W] layout.
"""
from __future__ import annotations

import ctypes as C
import warnings
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn

from .... import _lib
from ..mmdt import _no_grad_only

_warned_divisibility = False


@dataclass
class VideoVAEConfig:
    """Reference: vae_video3d.py:46-75."""
    in_ch: int = 3
    lat_ch: int = 8
    t_down: int = 4
    s_down: int = 8
    enc_base: int = 64
    enc_blocks: int = 2
    dec_base: int = 64
    dec_blocks: int = 2
    variational: bool = False
    out_activation: str = "sigmoid"

    @staticmethod
    def from_dict(d: Dict) -> "VideoVAEConfig":
        lat, enc, dec = d.get("latent", {}), d.get("encoder", {}), d.get("decoder", {})
        return VideoVAEConfig(in_ch=int(d.get("in_ch", 3)), lat_ch=int(lat.get("channels", 8)),
                              t_down=int(lat.get("t_down", 4)), s_down=int(lat.get("s_down", 8)),
                              enc_base=int(enc.get("base", 64)), enc_blocks=int(enc.get("blocks", 2)),
                              dec_base=int(dec.get("base", 64)), dec_blocks=int(dec.get("blocks", 2)),
                              variational=bool(d.get("variational", False)),
                              out_activation=str(d.get("out_activation", "sigmoid")))


def _block(c_in: int, c_out: int) -> nn.Sequential:
    """Conv3d(3x3x3, pad 1) -> GELU -> GroupNorm(min(8, C)) (vae_video3d.py:78-84)."""
    return nn.Sequential(nn.Conv3d(c_in, c_out, (3, 3, 3), padding=(1, 1, 1)), nn.GELU(),
                         nn.GroupNorm(num_groups=min(8, c_out), num_channels=c_out))


class VideoVAE(nn.Module):
    """Reference: vae_video3d.py:87-214."""

    def __init__(self, cfg: VideoVAEConfig):
        super().__init__()
        self.cfg = cfg
        if cfg.out_activation not in ("sigmoid", "tanh"):
            raise ValueError("out_activation must be 'sigmoid' or 'tanh'")
        c = cfg.enc_base
        enc = [_block(cfg.in_ch, c)] + [_block(c, c) for _ in range(cfg.enc_blocks - 1)]
        self.pool = nn.AvgPool3d(kernel_size=(cfg.t_down, cfg.s_down, cfg.s_down),
                                 stride=(cfg.t_down, cfg.s_down, cfg.s_down))
        self.enc_net = nn.Sequential(*enc)
        if cfg.variational:
            self.to_mu = nn.Conv3d(c, cfg.lat_ch, kernel_size=1)
            self.to_logv = nn.Conv3d(c, cfg.lat_ch, kernel_size=1)
        else:
            self.to_lat = nn.Conv3d(c, cfg.lat_ch, kernel_size=1)
        dch = cfg.dec_base
        self.from_lat = nn.Conv3d(cfg.lat_ch, dch, kernel_size=1)
        self.dec_net = nn.Sequential(*[_block(dch, dch) for _ in range(cfg.dec_blocks)])
        self.up_t, self.up_s = cfg.t_down, cfg.s_down
        self.to_img = nn.Conv3d(dch, cfg.in_ch, kernel_size=1)
        self.act_out = nn.Sigmoid() if cfg.out_activation == "sigmoid" else nn.Tanh()
        self._kld = None
        self._keep = []

    @classmethod
    def from_config(cls, d: Dict) -> "VideoVAE":
        return cls(VideoVAEConfig.from_dict(d))

    def kld_loss(self):
        return self._kld

    def _weights(self, device) -> _lib.VaeWeights:
        cfg = self.cfg
        if cfg.enc_base != 64 or cfg.dec_base != 64:
            raise NotImplementedError("the VideoVAE kernels are built for encoder.base = decoder.base = 64 (mvp.yaml)")
        if max(cfg.enc_blocks, cfg.dec_blocks) > _lib.AVD_VAE_MAX_BLOCKS or min(cfg.enc_blocks, cfg.dec_blocks) < 1:
            raise NotImplementedError(f"1..{_lib.AVD_VAE_MAX_BLOCKS} blocks per side")
        w = _lib.VaeWeights()
        w.in_ch, w.lat_ch, w.base, w.t_down, w.s_down = cfg.in_ch, cfg.lat_ch, 64, cfg.t_down, cfg.s_down
        w.enc_blocks, w.dec_blocks = cfg.enc_blocks, cfg.dec_blocks
        w.out_tanh = 1 if cfg.out_activation == "tanh" else 0
        keep = []

        def p(t: torch.Tensor) -> int:
            t = t.detach().to(device=device, dtype=torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        w.gn_eps = float(self.enc_net[0][2].eps)
        for i, blk in enumerate(self.enc_net):
            w.enc_w[i], w.enc_b[i], w.enc_gn_w[i], w.enc_gn_b[i] = p(blk[0].weight), p(blk[0].bias), p(blk[2].weight), p(blk[2].bias)
        for i, blk in enumerate(self.dec_net):
            w.dec_w[i], w.dec_b[i], w.dec_gn_w[i], w.dec_gn_b[i] = p(blk[0].weight), p(blk[0].bias), p(blk[2].weight), p(blk[2].bias)
        lat = self.to_mu if cfg.variational else self.to_lat          # eval: z = mu (vae_video3d.py:176-184)
        w.to_lat_w, w.to_lat_b = p(lat.weight), p(lat.bias)
        w.from_lat_w, w.from_lat_b = p(self.from_lat.weight), p(self.from_lat.bias)
        w.to_img_w, w.to_img_b = p(self.to_img.weight), p(self.to_img.bias)
        self._keep = keep
        return w

    def _check_divisible(self, T: int, H: int, W: int):
        """Reference: vae_video3d.py:139-162 -- center-crop to multiples of (t_down, s_down, s_down)."""
        global _warned_divisibility
        td, sd = self.cfg.t_down, self.cfg.s_down
        T2, H2, W2 = (T // td) * td, (H // sd) * sd, (W // sd) * sd
        if (T2, H2, W2) != (T, H, W) and not _warned_divisibility:
            warnings.warn(f"[VideoVAE] Input (T={T},H={H},W={W}) not divisible by (t_down={td}, s_down={sd}); "
                          f"center-cropping to (T={T2},H={H2},W={W2}).")
            _warned_divisibility = True
        t0, h0, w0 = (T - T2) // 2, (H - H2) // 2, (W - W2) // 2
        return T2, H2, W2, (t0, t0 + T2, h0, h0 + H2, w0, w0 + W2)

    def encode(self, x: torch.Tensor) -> torch.Tensor:
        _no_grad_only(self)
        if self.training and self.cfg.variational:
            raise NotImplementedError("the reparameterised (training-mode) encode is outside the sampling path")
        xc = _lib.require_cuda_f32(x, "x")
        if xc.dim() != 5 or xc.size(1) != self.cfg.in_ch:
            raise ValueError(f"VideoVAE.encode expects [B,{self.cfg.in_ch},T,H,W], got {tuple(xc.shape)}")
        B, _, T, H, W = xc.shape
        T2, H2, W2, (t0, t1, h0, h1, w0, w1) = self._check_divisible(T, H, W)
        if (T2, H2, W2) != (T, H, W):
            xc = xc[:, :, t0:t1, h0:h1, w0:w1].contiguous()
        if min(T2, H2, W2) == 0:
            raise ValueError("input smaller than one latent cell")
        lib = _lib.load()
        wts = self._weights(xc.device)
        z = torch.empty(B, self.cfg.lat_ch, T2 // self.cfg.t_down, H2 // self.cfg.s_down, W2 // self.cfg.s_down,
                        device=xc.device, dtype=torch.float32)
        scratch = torch.empty(int(lib.avd_vae_scratch_bytes(B, T2, H2, W2)), device=xc.device, dtype=torch.uint8)
        with torch.cuda.device(xc.device):
            _lib.check(lib.avd_video_encode(C.byref(wts), xc.data_ptr(), z.data_ptr(), scratch.data_ptr(), B, T2, H2, W2,
                                            _lib.stream_ptr(xc.device)))
        self._kld = None
        return z

    def decode(self, z: torch.Tensor, out_size: Optional[Tuple[int, int, int]] = None) -> torch.Tensor:
        _no_grad_only(self)
        zc = _lib.require_cuda_f32(z, "z")
        if zc.dim() != 5 or zc.size(1) != self.cfg.lat_ch:
            raise ValueError(f"VideoVAE.decode expects [B,{self.cfg.lat_ch},T',H',W'], got {tuple(zc.shape)}")
        B, _, Tl, Hl, Wl = zc.shape
        T, H, W = (Tl * self.up_t, Hl * self.up_s, Wl * self.up_s) if out_size is None else (int(s) for s in out_size)
        lib = _lib.load()
        wts = self._weights(zc.device)
        x = torch.empty(B, self.cfg.in_ch, T, H, W, device=zc.device, dtype=torch.float32)
        scratch = torch.empty(int(lib.avd_vae_scratch_bytes(B, T, H, W)), device=zc.device, dtype=torch.uint8)
        with torch.cuda.device(zc.device):
            _lib.check(lib.avd_video_decode(C.byref(wts), zc.data_ptr(), x.data_ptr(), scratch.data_ptr(), B, Tl, Hl, Wl,
                                            T, H, W, _lib.stream_ptr(zc.device)))
        return x
