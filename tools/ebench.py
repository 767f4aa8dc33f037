"""HBM roofline of the byte-moving kernels on one B200 (measurement tool; SURVEY.md §8(d) algorithmic bytes).

    python tools/ebench.py [--clips 2048] [--out gpurun_out/ebench.json]

At the mvp.yaml sampling batch (8 clips) these kernels move tens of kilobytes and last ~2 us, so their bandwidth there
is launch latency, not a statement about the kernel.  Here each one is run over `clips` clips (buffers several times
the 126 MB L2, so every launch streams from HBM), timed with CUDA events over back-to-back launches, and reported as
algorithmic bytes / time against MEASURED_PEAKS.json `hbm_gbs`:
  tube patch / unpatch      8 B per latent element (4 read + 4 written)
  audio chunk / unchunk     8 B per token element
  fused CFG + un-tokenise + DDIM (video, audio)   16 B per target-latent element (eps_c, eps_n, z_t read; z_prev written)
  ddim_step (standalone)    12 B per element
  rows_prepare (fp32 rows -> bf16 operand + row sums of squares)   6 B per element
  crossfade_audio / crossfade_video (stream stitching)   4 + 4 B per window / output sample; 1 + 1 B per pixel byte
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from multimodal_diffusion_b200 import _lib  # noqa: E402


def timed(fn, reps: int = 10) -> float:
    """microseconds per launch: `reps` back-to-back launches between two events, best of 3."""
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / reps)
    return best


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--clips", type=int, default=2048)
    ap.add_argument("--out", default="gpurun_out/ebench.json")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    st = _lib.stream_ptr
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    peak = float(peaks.get("hbm_gbs", 6460.2))
    B = args.clips
    rows = []

    def report(name, us, nbytes):
        gbs = nbytes / us * 1e-3
        rows.append({"kernel": name, "clips": B, "us": round(us, 1), "algorithmic_MB": round(nbytes / 1e6, 1),
                     "GBs": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 3)})
        print(json.dumps(rows[-1]), flush=True)

    zv = torch.randn(B, 8, 12, 16, 16, device=dev)
    tok = torch.empty(B, 96, 256, device=dev)
    report("tube_patch_video", timed(lambda: _lib.check(lib.avd_tube_patch(zv.data_ptr(), tok.data_ptr(), B, 8, 12, 16, 16, 2, 4, 4, st()))), zv.numel() * 8)
    zo = torch.empty_like(zv)
    report("tube_unpatch_video", timed(lambda: _lib.check(lib.avd_tube_unpatch(tok.data_ptr(), zo.data_ptr(), B, 8, 12, 16, 16, 2, 4, 4, st()))), zv.numel() * 8)

    Ba = B * 64                      # audio latents are 164x smaller per clip: more clips for the same bytes
    za = torch.randn(Ba, 8, 150, device=dev)
    ta = torch.empty(Ba, 37, 32, device=dev)
    report("chunk_1d (audio tokenise)", timed(lambda: _lib.check(lib.avd_chunk1d(za.data_ptr(), ta.data_ptr(), Ba, 8, 150, 4, 4, st()))), ta.numel() * 8)
    zao = torch.empty_like(za)
    report("overlap_add_1d (audio un-tokenise)", timed(lambda: _lib.check(lib.avd_unchunk1d(ta.data_ptr(), zao.data_ptr(), Ba, 8, 37, 4, 4, 150, st()))), ta.numel() * 8)

    ab = torch.linspace(0.9999, 1e-4, 1000, device=dev)
    ec, en = torch.randn(B, 96, 256, device=dev), torch.randn(B, 96, 256, device=dev)
    a = _lib.CfgDdimArgs(target=_lib.AVD_TARGET_VIDEO, B=B, C=8, T=12, H=16, W=16, tt=2, th=4, tw=4, guidance=3.5,
                         t_now=499, t_prev=479, T_train=1000)
    report("cfg + tube_unpatch + ddim (video)", timed(lambda: _lib.check(lib.avd_cfg_untok_ddim(C.byref(a), ec.data_ptr(), en.data_ptr(), zv.data_ptr(), zo.data_ptr(), ab.data_ptr(), st()))), zv.numel() * 16)
    eca, ena = torch.randn(Ba, 37, 32, device=dev), torch.randn(Ba, 37, 32, device=dev)
    aa = _lib.CfgDdimArgs(target=_lib.AVD_TARGET_AUDIO, B=Ba, C=8, L=150, length=4, stride=4, Nt=37, guidance=3.5,
                          t_now=499, t_prev=479, T_train=1000)
    report("cfg + overlap_add + ddim (audio)", timed(lambda: _lib.check(lib.avd_cfg_untok_ddim(C.byref(aa), eca.data_ptr(), ena.data_ptr(), za.data_ptr(), zao.data_ptr(), ab.data_ptr(), st()))), za.numel() * 16)

    tn = torch.full((B,), 499, device=dev, dtype=torch.int64)
    tp = torch.full((B,), 479, device=dev, dtype=torch.int64)
    per = zv.numel() // B
    report("ddim_step (standalone, eta=0)", timed(lambda: _lib.check(lib.avd_ddim_step(zv.data_ptr(), tok.data_ptr(), None, zo.data_ptr(), tn.data_ptr(), tp.data_ptr(), ab.data_ptr(), 1000, B, per, 0.0, st()))), zv.numel() * 12)

    M = B * 96
    x = torch.randn(M, 512, device=dev)
    xa = torch.empty(1, M, 512, device=dev, dtype=torch.bfloat16)
    ssq = torch.empty(M, 8, device=dev)
    report("rows_prepare (fp32 -> bf16 operand + ssq)", timed(lambda: _lib.check(lib.avd_rows_prepare(x.data_ptr(), xa.data_ptr(), ssq.data_ptr(), M, 512, _lib.AVD_PREC_BF16, st()))), x.numel() * 6)

    # streaming stitchers: every window sample read once, every output sample written once
    nw, Lw, hopw = 1024, 48000, 16000                       # 1024 windows of 3 s at 16 kHz, 1 s hop
    chunks = torch.randn(nw, Lw, device=dev)
    wnd = torch.ones(Lw, device=dev)
    outa = torch.empty((nw - 1) * hopw + Lw, device=dev)
    report("crossfade_audio (3 s windows, 1 s hop)", timed(lambda: _lib.check(lib.avd_crossfade_audio(chunks.data_ptr(), wnd.data_ptr(), outa.data_ptr(), nw, Lw, hopw, st()))), chunks.numel() * 4 + outa.numel() * 4)
    del chunks, outa
    nv, Tv, Fv, hopv = 192, 48, 128 * 128 * 3, 16           # 192 windows of 48 frames 128x128x3, 16-frame hop
    vch = torch.randint(0, 256, (nv, Tv, Fv), device=dev, dtype=torch.uint8)
    wv = torch.ones(Tv, device=dev)
    outv = torch.empty(((nv - 1) * hopv + Tv, Fv), device=dev, dtype=torch.uint8)
    report("crossfade_video (48-frame windows, 16-frame hop)", timed(lambda: _lib.check(lib.avd_crossfade_video(vch.data_ptr(), wv.data_ptr(), outv.data_ptr(), nv, Tv, Fv, hopv, st()))), vch.numel() + outv.numel())
    del vch, outv

    # encoders / decoders either side of the sampling loop: CUDA-core fp32 direct convolutions, run once per
    # clip.  Bound = the fp32 FMA rate (148 SMs x 128 lanes x 2 flop x SM clock), not HBM; the line to read is their time
    # next to one 50-step sampling call of the same 8 clips (16.3 ms).
    import numpy as np
    sys.path.insert(0, str(ROOT / "tests"))
    from multimodal_diffusion_b200.avdiff.models.infer import sample_clip as sc
    cfg = json.loads((ROOT / "tests" / "golden" / "cfg_mvp_v2a.json").read_text())
    torch.manual_seed(0)
    vae, codec = sc.build_components(cfg, dev)[:2]
    fp32_peak = 148 * 128 * 2 * float(peaks.get("sm_max_mhz", 1965.0)) * 1e-6     # TFLOP/s
    Be = 8
    frames = torch.rand(Be, 3, 48, 128, 128, device=dev)
    wav = torch.tanh(torch.randn(Be, 1, 48000, device=dev))
    enc_rows = []

    def conv_flops(mod, voxels_of):
        tot = 0.0
        for k, v in mod.state_dict().items():
            if v.dim() >= 3 and k.endswith("weight"):
                tot += 2.0 * v.numel() * voxels_of(k)
        return tot

    with torch.no_grad():
        zv8 = vae.encode(frames)
        za8 = codec.encode(wav)
        vox = 48 * 128 * 128
        lat = zv8.shape[2] * zv8.shape[3] * zv8.shape[4]
        cases = [
            ("VideoVAE.encode (8 clips, 3 s, 128x128)", lambda: vae.encode(frames),
             Be * conv_flops(vae, lambda k: vox if "enc_net" in k else (lat if "to_lat" in k else 0))),
            ("VideoVAE.decode", lambda: vae.decode(zv8),
             Be * conv_flops(vae, lambda k: vox if ("dec_net" in k or "to_img" in k) else (lat if "from_lat" in k else 0))),
            ("AudioCodec.encode (8 clips, 3 s, 16 kHz)", lambda: codec.encode(wav),
             Be * conv_flops(codec, lambda k: 48000 if "pre" in k else (za8.shape[-1] if "to_lat" in k else 0))),
            ("AudioCodec.decode", lambda: codec.decode(za8),
             Be * conv_flops(codec, lambda k: 48000 if "smooth" in k else (za8.shape[-1] if "from_lat" in k else 0))),
        ]
        for name, fn, fl in cases:
            us = timed(fn, reps=3)
            enc_rows.append({"kernel": name, "us": round(us, 1), "GFLOP": round(fl / 1e9, 2),
                             "TFLOPs_fp32": round(fl / us * 1e-6, 2), "frac_of_fp32_fma_peak": round(fl / us * 1e-6 / fp32_peak, 3)})
            print(json.dumps(enc_rows[-1]), flush=True)

    out = {"what": "byte-moving kernels at HBM-streaming sizes on 1 x B200; algorithmic bytes / CUDA-event time",
           "encoders": {"what": "encode / decode kernels (avd_codec.cu, avd_vae.cu) through the mirror modules, 8 clips; algorithmic "
                                "conv FLOPs (2 x weights x output positions) / CUDA-event time against the fp32 FMA peak",
                        "fp32_fma_peak_tflops": round(fp32_peak, 1), "rows": enc_rows},
           "hbm_peak_gbs": peak, "peak_source": "MEASURED_PEAKS.json hbm_gbs (torch copy, read+write bytes)", "rows": rows}
    Path(args.out).parent.mkdir(parents=True, exist_ok=True)
    Path(args.out).write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
