#!/usr/bin/env python3
"""Generate tests/golden/stream.npz by running the UNMODIFIED reference stitching functions on seeded inputs.

    PYTHONPATH=/root/reference python tests/golden/make_golden_stream.py

Build-container only (the reference does not travel to the GPU box).  The reference module imports
`avdiff.infer.sample_clip`, a package path that does not exist in its tree (the files live under
avdiff/models/infer/), so the script aliases `avdiff.infer` to `avdiff.models.infer` before importing it; the four
functions exercised here are then the reference's own code objects.
"""
from __future__ import annotations

import importlib
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, "/root/reference")
OUT = Path(__file__).resolve().parent

infer_pkg = importlib.import_module("avdiff.models.infer")
sys.modules["avdiff.infer"] = infer_pkg
sys.modules["avdiff.infer.sample_clip"] = importlib.import_module("avdiff.models.infer.sample_clip")
ref = importlib.import_module("avdiff.models.infer.stream_infer")


def main() -> None:
    rng = np.random.default_rng(20261104)
    d = {}
    # audio: sr 1000 keeps the fixture small; windows 3 s / hop 1 s / fade 0.25 s are mvp.yaml's streaming values
    y = rng.standard_normal(7300).astype(np.float32)
    ch, win, hop = ref.split_audio_into_windows(y, sr=1000, win_s=3.0, hop_s=1.0)
    d["a_y"], d["a_chunks"], d["a_win_hop"] = y, ch, np.array([win, hop])
    gen = rng.standard_normal(ch.shape).astype(np.float32)          # stands for the per-window generated audio
    d["a_gen"] = gen
    d["a_fade"] = ref.crossfade_audio(gen, sr=1000, hop=hop, win=win, fade_s=0.25)
    d["a_rect"] = ref.crossfade_audio(gen, sr=1000, hop=hop, win=win, fade_s=0.0)
    d["a_short"] = ref.split_audio_into_windows(y[:2500], sr=1000, win_s=3.0, hop_s=1.0)[0]
    # odd sizes: hop not a multiple of 4, hop > window (gaps), fade window halves overlapping
    gen2 = rng.standard_normal((5, 1001)).astype(np.float32)
    d["a_gen2"] = gen2
    d["a_odd"] = ref.crossfade_audio(gen2, sr=1000, hop=333, win=1001, fade_s=0.1)
    d["a_gap"] = ref.crossfade_audio(gen2, sr=1000, hop=1200, win=1001, fade_s=0.05)
    d["a_overlapfade"] = ref.crossfade_audio(gen2, sr=1000, hop=500, win=1001, fade_s=0.7)
    # video: 16 fps, 20x12 frames
    fr = rng.integers(0, 256, size=(101, 12, 20, 3), dtype=np.uint8)
    vch, vwin, vhop = ref.split_frames_into_windows(fr, fps=16, win_s=3.0, hop_s=1.0)
    d["v_frames"], d["v_chunks"], d["v_win_hop"] = fr, vch, np.array([vwin, vhop])
    vgen = rng.integers(0, 256, size=vch.shape, dtype=np.uint8)
    d["v_gen"] = vgen
    d["v_fade"] = ref.crossfade_video(vgen, hop=vhop, win=vwin, fade_f=4)
    d["v_nofade"] = ref.crossfade_video(vgen, hop=vhop, win=vwin, fade_f=0)
    vgen2 = rng.integers(0, 256, size=(3, 7, 5, 3, 3), dtype=np.uint8)      # 45-byte frames: the byte path
    d["v_gen2"] = vgen2
    d["v_odd"] = ref.crossfade_video(vgen2, hop=5, win=7, fade_f=3)
    d["v_short"] = ref.split_frames_into_windows(fr[:30], fps=16, win_s=3.0, hop_s=1.0)[0]
    # ops.overlap_add_1d(apply_hann=True) (avdiff/utils/ops.py:82-92): same arithmetic with a Hann window, clamp 1e-8
    import torch
    from avdiff.utils import ops
    hw = rng.standard_normal((5, 3, 9, 5)).astype(np.float32)
    d["hann_in"] = hw
    for s in (3, 5, 7):
        d[f"hann_s{s}"] = ops.overlap_add_1d(torch.from_numpy(hw), stride=s, apply_hann=True).numpy()
    hw2 = rng.standard_normal((2, 8, 37, 16)).astype(np.float32)
    d["hann_in2"] = hw2
    d["hann2_s4"] = ops.overlap_add_1d(torch.from_numpy(hw2), stride=4, apply_hann=True).numpy()
    np.savez_compressed(OUT / "stream.npz", **d)
    print({k: v.shape for k, v in d.items()})


if __name__ == "__main__":
    main()
