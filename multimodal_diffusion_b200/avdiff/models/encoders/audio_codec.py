"""audio_codec.py -- mirror of avdiff/models/encoders/audio_codec.py: waveform <-> latent frames.

Same constructor, parameter names (``pre.{0,1}.0``, ``to_lat``, ``from_lat``, ``smooth.{0,2,4}``), seeded init and
``encode`` / ``decode`` contracts as the reference, so its ``state_dict`` loads strictly; the arithmetic runs in
libavdiff_b200.so (csrc/avd_codec.cu: one kernel for encode, two for decode), fp32, rel-L2 ~1e-6 against the reference
module (tests/golden/encoders.npz).  There is no CPU path: CPU tensors raise.

    wav [B,1,L] --encode--> z [B,Ca,Fa]      Fa = frames_per_clip (exact) or ceil(L / hop_samples)
    z [B,Ca,Fa] --decode--> wav [B,1,Fa*hop_samples]
"""
from __future__ import annotations

import math
import warnings
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn

from .... import _lib
from ..mmdt import _no_grad_only


@dataclass
class AudioCodecConfig:
    """Reference: audio_codec.py:37-75."""
    in_ch: int = 1
    lat_ch: int = 8
    sr: int = 16000
    hop_samples: int = 320
    hidden: int = 64
    smooth_kernel: int = 7
    frames_per_clip: Optional[int] = None

    @staticmethod
    def from_dict(d: Dict) -> "AudioCodecConfig":
        lat, codec, sr = d.get("latent", {}), d.get("codec", {}), int(d.get("sr", 16000))
        hop = max(1, int(round(sr * float(lat["frame_hop_ms"]) / 1000.0))) if "frame_hop_ms" in lat \
            else int(codec.get("hop_samples", 320))
        return AudioCodecConfig(in_ch=int(d.get("in_ch", 1)), lat_ch=int(lat.get("channels", 8)), sr=sr, hop_samples=hop,
                                hidden=int(codec.get("hidden", 64)), smooth_kernel=int(codec.get("smooth_kernel", 7)),
                                frames_per_clip=int(lat.get("frames_per_clip", 0)) or None)


def _conv_act(c_in: int, c_out: int, k: int) -> nn.Sequential:
    return nn.Sequential(nn.Conv1d(c_in, c_out, kernel_size=k, padding=k // 2), nn.GELU())


class AudioCodec(nn.Module):
    """Reference: audio_codec.py:89-214."""

    def __init__(self, cfg: AudioCodecConfig):
        super().__init__()
        self.cfg = cfg
        k = max(3, int(cfg.smooth_kernel))
        self.pre = nn.Sequential(_conv_act(cfg.in_ch, cfg.hidden, 9), _conv_act(cfg.hidden, cfg.hidden, 9))
        self.to_lat = nn.Conv1d(cfg.hidden, cfg.lat_ch, kernel_size=1)
        self.from_lat = nn.Conv1d(cfg.lat_ch, cfg.hidden, kernel_size=1)
        self.smooth = nn.Sequential(nn.Conv1d(cfg.hidden, cfg.hidden, kernel_size=k, padding=k // 2), nn.GELU(),
                                    nn.Conv1d(cfg.hidden, cfg.hidden, kernel_size=k, padding=k // 2), nn.GELU(),
                                    nn.Conv1d(cfg.hidden, cfg.in_ch, kernel_size=k, padding=k // 2))
        for m in self.modules():
            if isinstance(m, nn.Conv1d):
                nn.init.kaiming_uniform_(m.weight, a=0.2)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)

    @classmethod
    def from_config(cls, d: Dict) -> "AudioCodec":
        return cls(AudioCodecConfig.from_dict(d))

    @property
    def hop(self) -> int:
        return int(self.cfg.hop_samples)

    @staticmethod
    def _compute_exact_pool_params(L: int, Fa: int) -> Tuple[int, int]:
        """Reference: audio_codec.py:137-147 -- integer hop with Fa * hop >= L."""
        assert Fa > 0
        hop = max(1, int(round(L / Fa)))
        if Fa * hop < L:
            hop += 1
        return hop, Fa * hop

    def _frames(self, L: int) -> Tuple[int, int]:
        """(Fa, hop) of _avgpool_frames (audio_codec.py:149-176)."""
        if self.cfg.frames_per_clip is None:
            return math.ceil(L / self.hop), self.hop
        Fa = int(self.cfg.frames_per_clip)
        return Fa, self._compute_exact_pool_params(L, Fa)[0]

    def _w(self, conv: nn.Conv1d):
        return conv.weight.detach().float().contiguous(), conv.bias.detach().float().contiguous()

    def encode(self, wav: torch.Tensor) -> torch.Tensor:
        _no_grad_only(self)
        assert wav.dim() == 3 and wav.size(1) == 1, "AudioCodec.encode expects [B,1,L]"
        if self.cfg.in_ch != 1:
            raise NotImplementedError("the encode kernel takes mono waveforms (in_ch = 1, as every config has)")
        x = _lib.require_cuda_f32(wav, "wav")
        B, _, L = x.shape
        Fa, hop = self._frames(L)
        z = torch.empty(B, self.cfg.lat_ch, Fa, device=x.device, dtype=torch.float32)
        (w0, b0), (w1, b1), (wl, bl) = self._w(self.pre[0][0]), self._w(self.pre[1][0]), self._w(self.to_lat)
        with torch.cuda.device(x.device):
            _lib.check(_lib.load().avd_audio_encode(x.data_ptr(), z.data_ptr(), B, L, self.cfg.hidden, self.cfg.lat_ch,
                                                    Fa, hop, w0.data_ptr(), b0.data_ptr(), w1.data_ptr(), b1.data_ptr(),
                                                    wl.data_ptr(), bl.data_ptr(), _lib.stream_ptr(x.device)))
        return z

    def decode(self, z: torch.Tensor) -> torch.Tensor:
        _no_grad_only(self)
        assert z.dim() == 3, "AudioCodec.decode expects [B,Ca,Fa]"
        if self.cfg.in_ch != 1:
            raise NotImplementedError("the decode kernel writes mono waveforms (in_ch = 1)")
        zc = _lib.require_cuda_f32(z, "z")
        B, Ca, Fa = zc.shape
        lib = _lib.load()
        k = int(self.smooth[0].kernel_size[0])
        wav = torch.empty(B, 1, Fa * self.hop, device=zc.device, dtype=torch.float32)
        scratch = torch.empty(int(lib.avd_audio_decode_scratch_bytes(B, self.cfg.hidden, Fa, self.hop)) // 4,
                              device=zc.device, dtype=torch.float32)
        (wf, bf), (w0, b0), (w2, b2), (w4, b4) = (self._w(m) for m in (self.from_lat, self.smooth[0], self.smooth[2],
                                                                        self.smooth[4]))
        with torch.cuda.device(zc.device):
            _lib.check(lib.avd_audio_decode(zc.data_ptr(), wav.data_ptr(), scratch.data_ptr(), B, Ca, Fa, self.cfg.hidden,
                                            self.hop, k, wf.data_ptr(), bf.data_ptr(), w0.data_ptr(), b0.data_ptr(),
                                            w2.data_ptr(), b2.data_ptr(), w4.data_ptr(), b4.data_ptr(),
                                            _lib.stream_ptr(zc.device)))
        return wav

    def check_consistency(self, clip_seconds: Optional[float] = None) -> None:
        """Reference: audio_codec.py:218-236."""
        if self.cfg.frames_per_clip is None or clip_seconds is None:
            return
        dur = self.cfg.frames_per_clip * self.hop / float(self.cfg.sr)
        if abs(dur - float(clip_seconds)) > 0.02:
            warnings.warn(f"[AudioCodec] frames_per_clip x hop = {dur:.3f} s but clips are {float(clip_seconds):.3f} s")
