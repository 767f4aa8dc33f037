"""stream_infer.py -- sliding-window driver around the fused sampler (mirror of avdiff/models/infer/stream_infer.py;
SURVEY.md §8(f) row N2).

Same names, argument meaning and return values as the reference for the four data functions:

    split_audio_into_windows(y, sr, win_s, hop_s)        reference :40-57   host index logic (numpy in, numpy out)
    split_frames_into_windows(frames, fps, win_s, hop_s) reference :60-79   host index logic
    crossfade_audio(chunks, sr, hop, win, fade_s)        reference :82-113  CUDA, avd_crossfade_audio, bit-exact
    crossfade_video(chunks, hop, win, fade_f)            reference :116-143 CUDA, avd_crossfade_video, bit-exact

The two stitchers run on the GPU: numpy arguments are copied to the device and the result comes back as numpy (what the
reference's callers expect); CUDA tensors are consumed in place and a CUDA tensor is returned.  The fade windows are
built on the host with numpy in fp32 exactly as the reference builds them, so their transcendental part cannot drift.

`stream_sample` replaces the reference's per-window Python loop (:182-188, :205-211): the windows of a long prompt are
independent, so they become the batch dimension of `FusedSampler` (and shard across ranks like clips do).  It works on
LATENT windows: the encoders / decoders either side (row N1) are not built yet.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import torch

from .... import _lib


# ------------------------------------------------------------------------------------------------ splitting (host)
def _window_starts(total: int, win: int, hop: int):
    """Start offsets the reference's while-loops visit (:47-56, :67-78): step by hop until a window reaches the end."""
    starts, s = [], 0
    while s < total:
        starts.append(s)
        if s + win >= total:
            break
        s += hop
    return starts


def split_audio_into_windows(y: np.ndarray, sr: int, win_s: float, hop_s: float) -> Tuple[np.ndarray, int, int]:
    """[L] -> ([n, win] zero-padded tail, win, hop).  A signal no longer than one window is returned whole and unpadded
    (reference :44-45)."""
    total = len(y)
    win, hop = int(round(sr * win_s)), int(round(sr * hop_s))
    if total <= win:
        return y[None, :], win, hop
    starts = _window_starts(total, win, hop)
    out = np.zeros((len(starts), win), dtype=y.dtype)
    for k, s in enumerate(starts):
        seg = y[s:s + win]
        out[k, :len(seg)] = seg
    return out, win, hop


def split_frames_into_windows(frames: np.ndarray, fps: int, win_s: float, hop_s: float) -> Tuple[np.ndarray, int, int]:
    """[T,H,W,3] -> ([n, win, H, W, 3] with the last frame repeated into a short tail, win, hop) (reference :60-79)."""
    total = frames.shape[0]
    win, hop = int(round(fps * win_s)), int(round(fps * hop_s))
    if total <= win:
        return frames[None, ...], win, hop
    starts = _window_starts(total, win, hop)
    out = np.empty((len(starts), win) + frames.shape[1:], dtype=frames.dtype)
    for k, s in enumerate(starts):
        seg = frames[s:s + win]
        out[k, :seg.shape[0]] = seg
        if seg.shape[0] < win:
            out[k, seg.shape[0]:] = seg[-1:]
    return out, win, hop


# ------------------------------------------------------------------------------------------------ fade windows (host)
def audio_fade_window(L: int, fade: int) -> np.ndarray:
    """fp32 [L]: ones with a raised-cosine fade-in over the first `fade` samples and fade-out over the last `fade`
    (reference :101-103); all ones for fade <= 0 (the rectangular overlap-add branch, :89-98).  Like the reference it
    raises ValueError when fade > L."""
    w = np.ones(L, dtype=np.float32)
    if fade > 0:
        phase = np.linspace(0, np.pi, fade, dtype=np.float32)
        w[:fade] = 0.5 * (1 - np.cos(phase))
        w[-fade:] = 0.5 * (1 + np.cos(phase))
    return w


def video_fade_window(L: int, fade: int) -> np.ndarray:
    """fp32 [L] per-frame weights: linear ramp 0 -> 1 over the first `fade` frames, 1 -> 0 over the last (reference
    :127-134).  The first and last weights are exactly 0, as in the reference."""
    w = np.ones(L, dtype=np.float32)
    if fade > 0:
        ramp = np.linspace(0, 1, fade, dtype=np.float32)
        w[:fade] *= ramp
        w[-fade:] *= ramp[::-1]
    return w


# ------------------------------------------------------------------------------------------------ stitching (CUDA)
def _to_device(a, dtype: torch.dtype) -> Tuple[torch.Tensor, bool]:
    if isinstance(a, torch.Tensor):
        if not a.is_cuda:
            raise ValueError("tensor arguments must live on the GPU (pass numpy arrays for host data)")
        return a.to(dtype).contiguous(), True
    return torch.from_numpy(np.ascontiguousarray(a)).to(dtype).cuda(non_blocking=False), False


def crossfade_audio(chunks, sr: int, hop: int, win: int, fade_s: float):
    """chunks [N, L] -> stitched [(N-1)*hop + L] fp32: overlap-add of the windows weighted by the fade window, divided by
    the summed weights (clamped at 1e-6).  Reference :82-113; `win` is unused there too."""
    if chunks.ndim != 2:
        raise ValueError("chunks must be [N, L]")
    N, L = int(chunks.shape[0]), int(chunks.shape[1])
    fade = int(round(sr * fade_s))
    w = audio_fade_window(L, fade)
    dev_chunks, on_gpu = _to_device(chunks, torch.float32)
    dev_w = torch.from_numpy(w).to(dev_chunks.device)
    out = torch.empty((N - 1) * hop + L, device=dev_chunks.device, dtype=torch.float32)
    _lib.check(_lib.load().avd_crossfade_audio(dev_chunks.data_ptr(), dev_w.data_ptr(), out.data_ptr(), N, L, int(hop),
                                               _lib.stream_ptr()))
    return out if on_gpu else out.cpu().numpy()


def crossfade_video(chunks, hop: int, win: int, fade_f: int):
    """chunks [N, T, H, W, 3] uint8 -> stitched frames [(N-1)*hop + T, H, W, 3] uint8 (alpha blend of overlapping
    frames with a triangular per-frame ramp).  Reference :116-143."""
    if chunks.ndim != 5:
        raise ValueError("chunks must be [N, T, H, W, C]")
    N, T, H, W, Cc = (int(v) for v in chunks.shape)
    w = video_fade_window(T, int(fade_f))
    dev_chunks, on_gpu = _to_device(chunks, torch.uint8)
    dev_w = torch.from_numpy(w).to(dev_chunks.device)
    out = torch.empty(((N - 1) * hop + T, H, W, Cc), device=dev_chunks.device, dtype=torch.uint8)
    _lib.check(_lib.load().avd_crossfade_video(dev_chunks.data_ptr(), dev_w.data_ptr(), out.data_ptr(), N, T,
                                               H * W * Cc, int(hop), _lib.stream_ptr()))
    return out if on_gpu else out.cpu().numpy()


# ------------------------------------------------------------------------------------------------ batched window loop
def stream_sample(sampler, z_prompt_windows: torch.Tensor, base_seed: int = 0,
                  z_init_windows: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Run the fused 50-step sampler over every window of a long prompt, `sampler.B` windows per call.

    z_prompt_windows [n, ...prompt latent shape] (CUDA fp32).  Initial noise of window k is `clip_noise(shape, k,
    base_seed)` unless `z_init_windows` is given, so the result does not depend on the batch size or on how the windows
    are sharded over ranks.  The last call is padded by repeating the final window; padded results are dropped.
    Replaces the reference's `for i in range(chunks.shape[0]): sample_one_direction(...)` loops (:182-188, :205-211)."""
    from ....sampler import clip_noise
    n = int(z_prompt_windows.shape[0])
    B = int(sampler.B)
    target_shape = tuple(sampler.latent_shape[1:])
    outs = []
    for s in range(0, n, B):
        idx = [min(s + k, n - 1) for k in range(B)]
        zp = z_prompt_windows[idx].contiguous()
        if z_init_windows is not None:
            zi = z_init_windows[idx].contiguous()
        else:
            zi = torch.stack([clip_noise(target_shape, i, base_seed) for i in idx]).to(zp.device)
        outs.append(sampler.sample(zp, zi)[: min(B, n - s)])
    return torch.cat(outs, dim=0)
