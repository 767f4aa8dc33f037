"""ops.py -- tokeniser index maps of the hot path on the GPU (mirror of avdiff/utils/ops.py).

Every function keeps the reference's name, signature, output shape and error behaviour; the data
movement is done by the bit-exact gather kernels of libavdiff_b200.so (K1 in DESIGN.md).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ... import _lib


def chunk_1d(x: torch.Tensor, length: int, stride: int, dim: int = -1) -> torch.Tensor:
    """Reference: avdiff/utils/ops.py:17-45.  windows[..., n, j] = x[..., n*stride + j],
    N = floor((L - length)/stride) + 1 (the tail is dropped).  C ABI: avd_chunk1d."""
    if dim < 0:
        dim += x.dim()
    transposed = dim != x.dim() - 1
    if transposed:
        x = x.transpose(dim, -1)
    L = x.size(-1)
    if length <= 0 or stride <= 0 or L < length:
        out = x[..., :max(0, min(L, length))].unsqueeze(-2)      # same single short window as the reference
    else:
        xc = _lib.require_cuda_f32(x, "x")
        lead = xc.shape[:-1]
        rows = int(xc.numel() // L) if L > 0 else 0
        n = (L - length) // stride + 1
        out = torch.empty(*lead, n, length, device=xc.device, dtype=torch.float32)
        if rows > 0:
            # [rows, 1, L] -> [rows, n, 1*length]
            _lib.check(_lib.load().avd_chunk1d(xc.data_ptr(), out.data_ptr(), rows, 1, L, length, stride,
                                               _lib.stream_ptr()))
    if transposed:
        out = out.transpose(dim, -1)
    return out


def overlap_add_1d(windows: torch.Tensor, stride: int, length: Optional[int] = None, dim_windows: int = -2,
                   apply_hann: bool = False) -> torch.Tensor:
    """Reference: avdiff/utils/ops.py:48-93.  Windows are summed in order i = 0..N-1 (the reference's
    sequential ``+=``) and divided by max(overlap count, 1e-8).  C ABI: avd_unchunk1d; with apply_hann=True the
    windows are weighted by torch.hann_window(W) and divided by the summed weights (avd_overlap_add_win)."""
    if length is None:
        length = windows.size(-1)
    nd = windows.dim()
    if dim_windows != nd - 2:
        perm = list(range(nd))
        perm.pop(dim_windows)
        perm.insert(nd - 2, dim_windows)
        windows = windows.permute(*perm)
    w = _lib.require_cuda_f32(windows, "windows")
    *prefix, N, W = w.shape
    L_out = (N - 1) * stride + W
    y = torch.empty(*prefix, L_out, device=w.device, dtype=torch.float32)
    rows = 1
    for s in prefix:
        rows *= int(s)
    if rows > 0 and L_out > 0 and apply_hann:
        win = torch.hann_window(W, dtype=torch.float32).to(w.device)     # the reference's window, computed by torch
        for r0 in range(0, rows, 65535):                                  # grid.y limit of the kernel
            r1 = min(rows, r0 + 65535)
            _lib.check(_lib.load().avd_overlap_add_win(w.reshape(rows, N, W)[r0:r1].data_ptr(), win.data_ptr(),
                                                       y.reshape(rows, L_out)[r0:r1].data_ptr(), r1 - r0, N, W,
                                                       stride, 1e-8, _lib.stream_ptr()))
    elif rows > 0 and L_out > 0:
        _lib.check(_lib.load().avd_unchunk1d(w.data_ptr(), y.data_ptr(), rows, 1, N, W, stride, L_out,
                                             _lib.stream_ptr()))
    return y


def tube_patch_video(z: torch.Tensor, t: int, h: int, w: int) -> torch.Tensor:
    """Reference: avdiff/utils/ops.py:100-119.  [B,C,T,H,W] -> [B,(T/t)(H/h)(W/w),C*t*h*w].  C ABI: avd_tube_patch."""
    B, C, T, H, W = z.shape
    assert T % t == 0 and H % h == 0 and W % w == 0, "tube sizes must divide latent dims"
    zc = _lib.require_cuda_f32(z, "z")
    N = (T // t) * (H // h) * (W // w)
    tokens = torch.empty(B, N, C * t * h * w, device=zc.device, dtype=torch.float32)
    if tokens.numel() > 0:
        _lib.check(_lib.load().avd_tube_patch(zc.data_ptr(), tokens.data_ptr(), B, C, T, H, W, t, h, w,
                                              _lib.stream_ptr()))
    return tokens


def tube_unpatch_video(tokens: torch.Tensor, C: int, T: int, H: int, W: int, t: int, h: int, w: int) -> torch.Tensor:
    """Reference: avdiff/utils/ops.py:122-144 (inverse of tube_patch_video).  C ABI: avd_tube_unpatch."""
    B, N, D = tokens.shape
    assert D == C * t * h * w, "token width mismatch"
    Tt, Hh, Ww = T // t, H // h, W // w
    assert N == Tt * Hh * Ww, "token count mismatch"
    tc = _lib.require_cuda_f32(tokens, "tokens")
    z = torch.empty(B, C, T, H, W, device=tc.device, dtype=torch.float32)
    if z.numel() > 0:
        _lib.check(_lib.load().avd_tube_unpatch(tc.data_ptr(), z.data_ptr(), B, C, T, H, W, t, h, w,
                                                _lib.stream_ptr()))
    return z


def pad_to_multiple(x: torch.Tensor, multiple: int, dim: int = -1, value: float = 0.0) -> Tuple[torch.Tensor, int]:
    """Reference: avdiff/utils/ops.py:151-161 (unused by the hot path; host-side allocation + copy)."""
    if dim < 0:
        dim += x.dim()
    size = x.size(dim)
    pad_amt = (multiple - size % multiple) % multiple
    if pad_amt == 0:
        return x, 0
    shape = list(x.shape)
    shape[dim] = size + pad_amt
    out = x.new_full(shape, value)
    out.narrow(dim, 0, size).copy_(x)
    return out, pad_amt


# ---- token-layout helpers used by the sampler and the tokenizer classes (sample_clip.py:179-215) ----

def latents_to_tokens_audio(z_a: torch.Tensor, l_chunk: int, s_chunk: int) -> torch.Tensor:
    """Reference: sample_clip.py:184-188 / tokenizers.py:83-88.  tok[b,n,c*l+j] = z[b,c,n*s+j] in one kernel."""
    zc = _lib.require_cuda_f32(z_a, "z_a")
    B, Ca, L = zc.shape
    if l_chunk <= 0 or s_chunk <= 0 or L < l_chunk:
        win = chunk_1d(zc, l_chunk, s_chunk, dim=-1)
        return win.permute(0, 2, 1, 3).contiguous().view(B, win.size(2), Ca * win.size(3))
    n = (L - l_chunk) // s_chunk + 1
    tok = torch.empty(B, n, Ca * l_chunk, device=zc.device, dtype=torch.float32)
    if tok.numel() > 0:
        _lib.check(_lib.load().avd_chunk1d(zc.data_ptr(), tok.data_ptr(), B, Ca, L, l_chunk, s_chunk,
                                           _lib.stream_ptr()))
    return tok


def tokens_to_latents_audio(tokens: torch.Tensor, Ca: int, l_chunk: int, Fa: int, stride: int) -> torch.Tensor:
    """Reference: sample_clip.py:191-215 / tokenizers.py:90-114: per-(b,c) overlap-add then crop / zero-pad to
    Fa, as ONE gather kernel (the reference runs B*Ca*Na Python-level slice updates)."""
    B, Na, D = tokens.shape
    assert D == Ca * l_chunk
    tc = _lib.require_cuda_f32(tokens, "tokens")
    z = torch.empty(B, Ca, Fa, device=tc.device, dtype=torch.float32)
    if z.numel() > 0:
        _lib.check(_lib.load().avd_unchunk1d(tc.data_ptr(), z.data_ptr(), B, Ca, Na, l_chunk, stride, Fa,
                                             _lib.stream_ptr()))
    return z
