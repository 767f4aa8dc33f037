"""GPU: avd_audio_decode_static -- the allocation-free variant of avd_audio_decode (scratch in a file-scope device
array) -- against avd_audio_decode with caller scratch, through the C-ABI.  Collected last on purpose (pytest -x).

Both entry points launch the same two kernels on the same inputs, so the outputs must be bit-identical; the check is
made on two calls in a row with different inputs (the second call finds the first call's data in the shared array) and
a third with a smaller shape, plus the size limit's error path."""
from __future__ import annotations

import pytest
import torch

from _common import randn

pytestmark = pytest.mark.gpu


def _weights(dev, H, Ca, K, seed):
    shapes = [(H, Ca, 1), (H,), (H, H, K), (H,), (H, H, K), (H,), (1, H, K), (1,)]
    scales = [Ca ** -0.5, 0.1, (H * K) ** -0.5, 0.1, (H * K) ** -0.5, 0.1, (H * K) ** -0.5, 0.1]
    return [(randn(seed + i, *s) * sc).to(dev).contiguous() for i, (s, sc) in enumerate(zip(shapes, scales))]


def _decode(lib, _lib, dev, z, ws, H, hop, K, static):
    B, Ca, Fa = z.shape
    wav = torch.full((B, 1, Fa * hop), float("nan"), device=dev, dtype=torch.float32)
    wp = [w.data_ptr() for w in ws]
    with torch.cuda.device(dev):
        st = _lib.stream_ptr(dev)
        if static:
            rc = lib.avd_audio_decode_static(z.data_ptr(), wav.data_ptr(), B, Ca, Fa, H, hop, K, *wp, st)
        else:
            scratch = torch.empty(int(lib.avd_audio_decode_scratch_bytes(B, H, Fa, hop)) // 4, device=dev,
                                  dtype=torch.float32)
            rc = lib.avd_audio_decode(z.data_ptr(), wav.data_ptr(), scratch.data_ptr(), B, Ca, Fa, H, hop, K, *wp, st)
        _lib.check(rc)
        torch.cuda.synchronize(dev)
    return wav


def test_audio_decode_static_matches_caller_scratch(lib):
    from multimodal_diffusion_b200 import _lib
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    H, Ca, K = 64, 8, 7
    ws = _weights(dev, H, Ca, K, 900)
    # two calls in a row at the same shape with different latents, then a smaller shape with another hop
    cases = [(2, 37, 320, 11), (2, 37, 320, 12), (1, 9, 160, 13)]
    for B, Fa, hop, seed in cases:
        z = randn(seed, B, Ca, Fa).to(dev)
        want = _decode(lib, _lib, dev, z, ws, H, hop, K, static=False)
        got = _decode(lib, _lib, dev, z, ws, H, hop, K, static=True)
        assert torch.isfinite(want).all() and want.abs().max() > 1e-3
        assert torch.equal(got, want), f"static-scratch decode differs at B={B} Fa={Fa} hop={hop}"
    # the limit: 16 MiB of scratch = B*H*Fa*hop*4
    limit = _lib.AVD_AUDIO_DECODE_STATIC_MAX_BYTES
    B, Fa, hop = 1, 205, 320                                   # 64 * 65600 * 4 = 16.8 MB > 16 MiB
    assert B * H * Fa * hop * 4 > limit
    z = randn(14, B, Ca, Fa).to(dev)
    wav = torch.zeros(B, 1, Fa * hop, device=dev)
    wp = [w.data_ptr() for w in ws]
    with torch.cuda.device(dev):
        rc = lib.avd_audio_decode_static(z.data_ptr(), wav.data_ptr(), B, Ca, Fa, H, hop, K, *wp, _lib.stream_ptr(dev))
    assert rc < 0
    with pytest.raises(_lib.AvdError, match="exceed the built-in"):
        _lib.check(rc)
    torch.cuda.synchronize(dev)
    assert float(wav.abs().max()) == 0.0                       # nothing was launched
    # the largest size that fits still runs (204 frames of 320: 16.7 MB <= 16 MiB)
    Fa = 204
    assert B * H * Fa * hop * 4 <= limit
    z = randn(15, B, Ca, Fa).to(dev)
    assert torch.equal(_decode(lib, _lib, dev, z, ws, H, hop, K, static=True),
                       _decode(lib, _lib, dev, z, ws, H, hop, K, static=False))
