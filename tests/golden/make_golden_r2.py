#!/usr/bin/env python3
"""Round-2 fixtures from the UNMODIFIED reference (CPU fp32):

    PYTHONPATH=/root/reference python tests/golden/make_golden_r2.py

  embeddings.npz     adapters.ModalityEmbedding / PositionalEmbedding1D / PositionalEmbedding3D / TimestepEmbedder
                     (adapters.py:41-158) on seeded parameters and inputs -- SURVEY §8 a14 / N4(iii)
  head_variants.npz  MultiModalNoiseHead with leaky_relu / relu activations, modality-specific trunks and share_parameters
                     (noise_heads.py:110-229) -- N4(iv)
  trainer_path.npz   the trainer's input path (trainer.py:38-49,195-202: full-width adapters + ADDITIVE sinusoidal
                     embedding) and one reference forward through it -- N3
  encoders.npz       AudioCodec / VideoVAE encode + decode (audio_codec.py:184-214, vae_video3d.py:164-214) on small
                     seeded inputs -- N1

Weights are re-created from seeds by the tests (same constructor sequence, checked by checksums)."""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference")
sys.path.insert(0, str(REF))
OUT = Path(__file__).resolve().parent

from avdiff.models import adapters as ad                                  # noqa: E402
from avdiff.models.heads.noise_heads import MultiModalNoiseHead           # noqa: E402
from avdiff.models.mmdt import MMDiT                                      # noqa: E402
from avdiff.models.train import trainer as tr                             # noqa: E402
from avdiff.models.encoders.audio_codec import AudioCodec, AudioCodecConfig   # noqa: E402
from avdiff.models.encoders.vae_video3d import VideoVAE, VideoVAEConfig   # noqa: E402
from avdiff.utils import ops                                              # noqa: E402


def randn(seed, *shape):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed), dtype=torch.float32)


def checksums(mod):
    keys, vals = [], []
    for k, v in mod.state_dict().items():
        keys.append(k)
        vals.append([float(v.double().sum()), float(v.double().abs().sum())])
    return np.array(keys), np.array(vals, dtype=np.float64)


def golden_embeddings():
    g = {}
    x = randn(1, 2, 37, 256)
    torch.manual_seed(200); me = ad.ModalityEmbedding(256)
    g["mod_x"] = x.numpy()
    g["mod_video"] = me(x, "video").detach().numpy()
    g["mod_audio"] = me(x, "audio").detach().numpy()
    torch.manual_seed(201); p1 = ad.PositionalEmbedding1D(256, max_len=64, mode="learned")
    g["pos1d_learned"] = p1(2, 37, torch.device("cpu"), torch.float32).detach().numpy()
    g["pos1d_sin"] = ad.PositionalEmbedding1D(256, mode="sin")(2, 37, torch.device("cpu"), torch.float32).numpy()
    torch.manual_seed(202); p3 = ad.PositionalEmbedding3D(256, max_T=8, max_H=8, max_W=8, mode="learned")
    g["pos3d_learned"] = p3(2, 6, 4, 4, torch.device("cpu"), torch.float32).detach().numpy()
    g["pos3d_sin"] = ad.PositionalEmbedding3D(256, mode="sin")(2, 6, 4, 4, torch.device("cpu"), torch.float32).numpy()
    t = torch.tensor([0, 19, 500, 999])
    g["t"] = t.numpy()
    g["temb_sin"] = ad.TimestepEmbedder(ad.TimestepCfg(dim=256, mode="sin"))(t).numpy()
    torch.manual_seed(203); te = ad.TimestepEmbedder(ad.TimestepCfg(dim=64, mode="mlp"))
    g["temb_mlp"] = te(t).detach().numpy()
    np.savez_compressed(OUT / "embeddings.npz", **g)
    print("embeddings.npz", {k: v.shape for k, v in g.items()})


def golden_head_variants():
    g = {}
    h = randn(5, 2, 9, 128)
    g["h"] = h.numpy()
    cases = {"leaky_spec2": dict(activation="leaky_relu", num_shared_layers=1, num_modality_specific_layers=3),
             "relu_shared_spec": dict(activation="relu", num_shared_layers=2, num_modality_specific_layers=2,
                                      share_parameters=True),
             "gelu_noshared": dict(activation="gelu", num_shared_layers=0, num_modality_specific_layers=2)}
    for i, (name, kw) in enumerate(cases.items()):
        torch.manual_seed(300 + i)
        head = MultiModalNoiseHead(input_dims={"video": 128, "audio": 128}, output_dims={"video": 64, "audio": 32},
                                   hidden_dim=128, dropout=0.0, **kw).eval()
        with torch.no_grad():                      # non-trivial LayerNorm affine parameters
            for k, v in head.state_dict().items():
                if k.endswith(".1.weight"):
                    v.copy_(1.0 + 0.2 * randn(len(k), *v.shape))
                elif k.endswith(".1.bias"):
                    v.copy_(0.1 * randn(len(k) + 1, *v.shape))
            out = head({"video": h, "audio": h})
        g[f"{name}_video"] = out["video"].numpy()
        g[f"{name}_audio"] = out["audio"].numpy()
    np.savez_compressed(OUT / "head_variants.npz", **g)
    print("head_variants.npz", {k: v.shape for k, v in g.items()})


def golden_trainer_path():
    """trainer.py:195-202 wiring with the mvp core: X = adapt(tok) + emb(t) (additive, width d)."""
    g = {}
    torch.manual_seed(400); adapt_v = tr.LinearAdapter(256, 512)
    torch.manual_seed(401); adapt_a = tr.LinearAdapter(32, 512)
    z_v, z_a = randn(11, 1, 8, 12, 16, 16), randn(12, 1, 8, 150)
    tok_v = ops.tube_patch_video(z_v, t=2, h=4, w=4)
    win = ops.chunk_1d(z_a, length=4, stride=4, dim=-1)
    tok_a = win.permute(0, 2, 1, 3).contiguous().view(1, win.shape[2], -1)
    t_v, t_a = torch.tensor([0]), torch.tensor([979])
    with torch.no_grad():
        Xv = tr.add_sinusoidal_timestep(adapt_v(tok_v), t_v, 256)       # `dim` is ignored by the trainer's variant
        Xa = tr.add_sinusoidal_timestep(adapt_a(tok_a), t_a, 256)
    g["z_v"], g["z_a"] = z_v.numpy(), z_a.numpy()
    g["X"] = torch.cat([Xv, Xa], dim=1).numpy()
    kv, cv = checksums(adapt_v)
    g["adapt_v_keys"], g["adapt_v_sums"] = kv, cv
    np.savez_compressed(OUT / "trainer_path.npz", **g)
    print("trainer_path.npz", {k: v.shape for k, v in g.items()})


def golden_encoders():
    g = {}
    torch.manual_seed(500)
    codec = AudioCodec(AudioCodecConfig(lat_ch=8, hop_samples=320, hidden=64, smooth_kernel=7, frames_per_clip=10)).eval()
    wav = torch.tanh(randn(21, 2, 1, 3200 - 37))                 # not a multiple of the hop: exercises the right pad
    with torch.no_grad():
        z = codec.encode(wav)
        wav_hat = codec.decode(z)
    g["wav"], g["z_a"], g["wav_hat"] = wav.numpy(), z.numpy(), wav_hat.numpy()
    torch.manual_seed(501)
    codec2 = AudioCodec(AudioCodecConfig(lat_ch=8, hop_samples=160, hidden=64, smooth_kernel=7, frames_per_clip=None)).eval()
    wav2 = torch.tanh(randn(22, 1, 1, 1700))
    with torch.no_grad():
        z2 = codec2.encode(wav2)
        g["wav2"], g["z_a2"], g["wav_hat2"] = wav2.numpy(), z2.numpy(), codec2.decode(z2).numpy()
    torch.manual_seed(502)
    vae = VideoVAE(VideoVAEConfig(lat_ch=8, t_down=4, s_down=8, enc_base=64, enc_blocks=2, dec_base=64, dec_blocks=2)).eval()
    with torch.no_grad():
        for k, v in vae.state_dict().items():                    # non-trivial GroupNorm affine parameters
            if ".2.weight" in k:
                v.copy_(1.0 + 0.2 * randn(len(k), *v.shape))
            elif ".2.bias" in k:
                v.copy_(0.1 * randn(len(k) + 1, *v.shape))
        x = torch.rand(1, 3, 8, 32, 32, generator=torch.Generator().manual_seed(23))
        zv = vae.encode(x)
        x_hat = vae.decode(zv)
    g["frames"], g["z_v"], g["frames_hat"] = x.numpy(), zv.numpy(), x_hat.numpy()
    kv, cv = checksums(vae)
    g["vae_keys"], g["vae_sums"] = kv, cv
    kc, cc = checksums(codec)
    g["codec_keys"], g["codec_sums"] = kc, cc
    np.savez_compressed(OUT / "encoders.npz", **g)
    print("encoders.npz", {k: v.shape for k, v in g.items()})


def golden_configs():
    """The merged YAML dicts sample_one_direction reads (load_config(mvp.yaml, v2a.yaml / a2v.yaml)), with the
    benchmark's 50 sampler steps (as tests/golden/make_golden.py drove the reference), as JSON."""
    import json
    from avdiff.utils.io import load_config
    for name, key in (("v2a", "audio"), ("a2v", "video")):
        cfg = load_config(str(REF / "configs/mvp.yaml"), str(REF / f"configs/{name}.yaml"))
        cfg["diffusion"][key]["sampler_steps"] = 50
        (OUT / f"cfg_mvp_{name}.json").write_text(json.dumps(cfg, indent=1, sort_keys=True, default=str))
        print(f"cfg_mvp_{name}.json", len(cfg))


if __name__ == "__main__":
    torch.set_num_threads(8)
    golden_configs()
    golden_embeddings()
    golden_head_variants()
    golden_trainer_path()
    golden_encoders()
