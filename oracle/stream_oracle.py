"""
stream_oracle.py -- CPU restatement (numpy, fp32) of the stitching functions of the reference's sliding-window driver.

TEST INFRASTRUCTURE ONLY (same rule as avdiff_oracle.py: only tests/, __graft_entry__.smoke() and bench.py's CPU legs
may import anything under oracle/; the product path never does).

Restates avdiff/models/infer/stream_infer.py of mauruszach/multimodal_diffusion:
    split_audio_into_windows   :40-57        split_frames_into_windows  :60-79
    crossfade_audio            :82-113       crossfade_video            :116-143
in scatter form -- window after window added into an output buffer and a weight buffer, then one divide -- which is the
order of fp32 additions the CUDA gather kernels (csrc/avd_stream.cu) have to reproduce bit for bit.

Parity pin: tests/golden/stream.npz holds outputs of the reference's own functions on seeded inputs
(tests/golden/make_golden_stream.py, run in the build container); tests/test_oracle_golden.py checks this file
against them bit-exactly.
"""
from __future__ import annotations

import numpy as np


def window_starts(total: int, win: int, hop: int):
    """offsets visited by the while-loops at :47-56 / :67-78"""
    out, s = [], 0
    while s < total:
        out.append(s)
        if min(total, s + win) == total:
            break
        s += hop
    return out


def split_audio(y: np.ndarray, sr: int, win_s: float, hop_s: float):
    win, hop = int(round(sr * win_s)), int(round(sr * hop_s))
    if len(y) <= win:                                   # :44-45
        return y[None, :], win, hop
    rows = []
    for s in window_starts(len(y), win, hop):
        seg = y[s:s + win]
        rows.append(np.concatenate([seg, np.zeros(win - len(seg), dtype=y.dtype)]))     # zero pad, :51-52
    return np.stack(rows), win, hop


def split_frames(frames: np.ndarray, fps: int, win_s: float, hop_s: float):
    win, hop = int(round(fps * win_s)), int(round(fps * hop_s))
    if frames.shape[0] <= win:                          # :64-65
        return frames[None, ...], win, hop
    rows = []
    for s in window_starts(frames.shape[0], win, hop):
        seg = frames[s:s + win]
        if seg.shape[0] < win:                          # repeat the last frame, :71-74
            seg = np.concatenate([seg, np.repeat(seg[-1:], win - seg.shape[0], axis=0)])
        rows.append(seg)
    return np.stack(rows), win, hop


def audio_window(L: int, fade: int) -> np.ndarray:
    w = np.ones(L, dtype=np.float32)
    if fade > 0:                                        # :101-103
        ph = np.linspace(0, np.pi, fade, dtype=np.float32)
        w[:fade] = 0.5 * (1 - np.cos(ph))
        w[-fade:] = 0.5 * (1 + np.cos(ph))
    return w


def video_window(L: int, fade: int) -> np.ndarray:
    w = np.ones(L, dtype=np.float32)
    if fade > 0:                                        # :127-134
        ramp = np.linspace(0, 1, fade, dtype=np.float32)
        w[:fade] *= ramp
        w[-fade:] *= ramp[::-1]
    return w


def crossfade_audio(chunks: np.ndarray, sr: int, hop: int, win: int, fade_s: float) -> np.ndarray:
    n, L = chunks.shape
    w = audio_window(L, int(round(sr * fade_s)))
    y = np.zeros((n - 1) * hop + L, dtype=np.float32)
    norm = np.zeros_like(y)
    for i in range(n):                                  # :106-110 (and :92-96 with w == 1)
        y[i * hop:i * hop + L] += chunks[i].astype(np.float32) * w
        norm[i * hop:i * hop + L] += w
    return (y / np.maximum(norm, np.float32(1e-6))).astype(np.float32)


def crossfade_video(chunks: np.ndarray, hop: int, win: int, fade_f: int) -> np.ndarray:
    n, L = chunks.shape[:2]
    w = video_window(L, int(fade_f)).reshape(L, 1, 1, 1)
    out = np.zeros(((n - 1) * hop + L,) + chunks.shape[2:], dtype=np.float32)
    norm = np.zeros(((n - 1) * hop + L, 1, 1, 1), dtype=np.float32)
    for i in range(n):                                  # :136-140
        out[i * hop:i * hop + L] += (chunks[i].astype(np.float32) / np.float32(255.0)) * w
        norm[i * hop:i * hop + L] += w
    out = out / np.maximum(norm, np.float32(1e-6))
    return (np.clip(out, 0, 1) * np.float32(255.0)).astype(np.uint8)    # :142-143
