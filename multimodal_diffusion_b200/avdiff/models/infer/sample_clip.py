"""sample_clip.py -- DDIM sampling with classifier-free guidance, V->A or A->V (mirror of
avdiff/models/infer/sample_clip.py's sampling surface; the CLI / file IO of the reference is out of scope).

``sample_one_direction`` keeps the reference's keyword signature and return value.  The encoders/decoders
(VideoVAE / AudioCodec) are the mirror modules of ``avdiff/models/encoders`` (CUDA kernels of
csrc/avd_codec.cu / avd_vae.cu); any other objects with ``encode`` / ``decode`` work too (stand-ins that inject latents
in the tests).  The loop
itself runs as the fused, CUDA-graph-captured kernel sequence of ``FusedSampler`` -- one sequence for the
cond and null CFG branches together -- instead of the reference's ~400 eager launches per step.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from ....sampler import DEFAULT_CORE_PRECISION, DEFAULT_HEAD_PRECISION, FusedSampler, HotPathConfig
from ...utils import ops
from ...utils import schedule_utils as su
from ..adapters import LinearAdapter as _XavierLinearAdapter
from ..heads.noise_heads import MultiModalNoiseHead
from ..mmdt import MMDiT


class LinearAdapter(_XavierLinearAdapter):
    """Reference: sample_clip.py:48-56 -- nn.Linear default init (the adapters.py variant uses xavier)."""

    def __init__(self, d_in: int, d_out: int):
        super().__init__(d_in, d_out, xavier=False)


def add_sinusoidal_timestep(tokens: torch.Tensor, t_scalar: torch.Tensor, dim: int) -> torch.Tensor:
    """Reference: sample_clip.py:59-70 -- CONCAT a broadcast sinusoidal embedding: [B,N,d] -> [B,N,d+dim]."""
    emb = su.timestep_embedding(t_scalar.to(tokens.device), dim=dim)
    return torch.cat([tokens, emb.unsqueeze(1).expand(-1, tokens.size(1), -1)], dim=-1)


def build_components(cfg: Dict, device: torch.device, vid_vae=None, aud_codec=None):
    """Reference: sample_clip.py:75-109.  Returns the same 7-tuple, built in the reference's order (VideoVAE, AudioCodec,
    adapters, core, head) so a seeded construction draws the same parameters.  vid_vae / aud_codec (extension): use
    these objects instead of constructing the mirror encoders (anything with ``encode`` / ``decode``)."""
    from ..encoders.audio_codec import AudioCodec
    from ..encoders.vae_video3d import VideoVAE
    if vid_vae is None:
        vid_vae = VideoVAE.from_config(cfg["video"]).to(device).eval()
    if aud_codec is None:
        aud_codec = AudioCodec.from_config(cfg["audio"]).to(device).eval()
    d = int(cfg["tokenizer"]["width"])
    out_v = int(cfg["model"]["heads"]["video"]["out_dim"])
    out_a = int(cfg["model"]["heads"]["audio"]["out_dim"])
    tstep_dim = int(cfg["embeddings"].get("timestep_dim", 256))
    adapt_v = LinearAdapter(out_v, d - tstep_dim).to(device)
    adapt_a = LinearAdapter(out_a, d - tstep_dim).to(device)
    core = MMDiT(**cfg["model"]["core"]).to(device).eval()
    head = MultiModalNoiseHead(
        input_dims={"video": d, "audio": d}, output_dims={"video": out_v, "audio": out_a},
        hidden_dim=int(cfg["model"]["heads"]["video"]["hidden_dim"]), num_shared_layers=2,
        num_modality_specific_layers=1, dropout=float(cfg["model"]["core"].get("dropout", 0.1)),
        activation=cfg["model"]["heads"]["video"].get("activation", "gelu")).to(device).eval()
    return vid_vae, aud_codec, adapt_v, adapt_a, core, head, tstep_dim


def latents_to_tokens_video(z_v: torch.Tensor, t_p: int, p: int) -> torch.Tensor:
    """Reference: sample_clip.py:179-181."""
    return ops.tube_patch_video(z_v, t=t_p, h=p, w=p)


def latents_to_tokens_audio(z_a: torch.Tensor, l_chunk: int, s_chunk: int) -> torch.Tensor:
    """Reference: sample_clip.py:184-188."""
    return ops.latents_to_tokens_audio(z_a, l_chunk, s_chunk)


def tokens_to_latents_audio(tokens: torch.Tensor, Ca: int, l_chunk: int, Fa: int, stride: int) -> torch.Tensor:
    """Reference: sample_clip.py:191-215."""
    return ops.tokens_to_latents_audio(tokens, Ca, l_chunk, Fa, stride)


# FusedSampler objects (packed weights + captured 50-step graph) are expensive to build: they are cached per
# (modules and their parameter versions, shapes, precision, device) and reused by later calls with the same modules.
_SAMPLER_CACHE: Dict[tuple, FusedSampler] = {}
_SAMPLER_CACHE_MAX = 8


def _module_key(m) -> tuple:
    if isinstance(m, dict):
        return tuple((k, v.data_ptr(), v._version) for k, v in m.items())
    return (id(m), bool(m.training)) + tuple((p.data_ptr(), p._version) for p in m.parameters())


def cached_sampler(hp: HotPathConfig, adapt_v, adapt_a, core, head, target: str, batch: int, device,
                   core_precision: str, head_precision: str, **kw) -> FusedSampler:
    import dataclasses
    key = (dataclasses.astuple(hp), target, int(batch), str(torch.device(device)), core_precision, head_precision,
           tuple(sorted((k, v if not isinstance(v, torch.Tensor) else (v.data_ptr(), v._version)) for k, v in kw.items())),
           _module_key(adapt_v), _module_key(adapt_a), _module_key(core), _module_key(head))
    fs = _SAMPLER_CACHE.get(key)
    if fs is None:
        if len(_SAMPLER_CACHE) >= _SAMPLER_CACHE_MAX:
            _SAMPLER_CACHE.pop(next(iter(_SAMPLER_CACHE)))
        fs = FusedSampler(hp, adapt_v, adapt_a, core, head, target, batch, device=device,
                          core_precision=core_precision, head_precision=head_precision, **kw)
        _SAMPLER_CACHE[key] = fs
    return fs


@torch.no_grad()
def sample_latents(*, cfg: Dict, adapt_v, adapt_a, core, head, target: str, z_prompt: torch.Tensor,
                   z_init: torch.Tensor, sampler_steps: Optional[int] = None,
                   core_precision: str = DEFAULT_CORE_PRECISION, head_precision: str = DEFAULT_HEAD_PRECISION,
                   use_graph: bool = True, seed: int = 0, clip_index0: int = 0) -> torch.Tensor:
    """Batched latent-space entry: the reference loop body (sample_clip.py:317-348 / 359-389) for B clips.
    z_prompt / z_init: CUDA fp32 latents; returns the final target latent.  Default precision = the timed and
    validated f16/f16 mode (sampler.DEFAULT_*_PRECISION); the FusedSampler behind it is cached across calls."""
    hp = HotPathConfig.from_yaml_dict(cfg, target, sampler_steps)
    if target == "video":
        hp.video_latent = tuple(int(s) for s in z_init.shape[2:])
        hp.Fa = int(z_prompt.shape[-1])
    else:
        hp.video_latent = tuple(int(s) for s in z_prompt.shape[2:])
        hp.Fa = int(z_init.shape[-1])
    kw = {"seed": int(seed), "clip_index0": int(clip_index0)} if hp.eta > 0.0 else {}
    fs = cached_sampler(hp, adapt_v, adapt_a, core, head, target, z_init.shape[0], z_init.device, core_precision,
                        head_precision, **kw)
    return fs.sample(z_prompt, z_init) if use_graph else fs.sample_eager(z_prompt, z_init)


@torch.no_grad()
def sample_one_direction(*, cfg: Dict, vid_vae, aud_codec, adapt_v, adapt_a, core, head, tstep_dim: int,
                         prompt_modality: str, prompt_video: Optional[np.ndarray], prompt_audio: Optional[np.ndarray],
                         device: torch.device, core_precision: str = DEFAULT_CORE_PRECISION,
                         head_precision: str = DEFAULT_HEAD_PRECISION, z_init: Optional[torch.Tensor] = None
                         ) -> Dict[str, np.ndarray]:
    """Reference: sample_clip.py:220-394 (B = 1).  Encode the prompt with the supplied encoder, run the fused
    DDIM+CFG loop on the GPU, decode with the supplied decoder.
    z_init (extension): the target's initial noise; by default it is drawn with torch.randn on `device` as the reference
    does (sample_clip.py:314 / 357) -- pass it to reproduce a run whose noise came from another generator."""
    if int(cfg["embeddings"].get("timestep_dim", 256)) != int(tstep_dim):
        raise ValueError("tstep_dim does not match cfg['embeddings']['timestep_dim']")
    Cv = int(cfg["video"]["latent"]["channels"])
    Ca = int(cfg["audio"]["latent"]["channels"])
    Fa = int(cfg["audio"]["latent"]["frames_per_clip"])
    fps, sr = int(cfg["video"]["fps"]), int(cfg["audio"]["sr"])
    B = 1
    if prompt_modality == "video":
        if prompt_video is None:
            raise ValueError("prompt_video frames required for prompt_modality=video")
        frames = torch.from_numpy(prompt_video).to(device).float() / 255.0            # [T,H,W,3]
        # the reference permutes to [1,T,3,H,W] here and its VideoVAE.encode then fails;
        # the layout the encoder documents is [1,3,T,H,W]:
        z_prompt = vid_vae.encode(frames.permute(3, 0, 1, 2).unsqueeze(0))
        z_init = torch.randn(B, Ca, Fa, device=device) if z_init is None else z_init.to(device).float()
        target = "audio"
    else:
        if prompt_audio is None:
            raise ValueError("prompt_audio required for prompt_modality=audio")
        z_prompt = aud_codec.encode(torch.from_numpy(prompt_audio).to(device).view(1, 1, -1))
        T_in = prompt_video.shape[0] if prompt_video is not None else int(round(cfg["data"]["clip_seconds"] * fps))
        H, W = int(cfg["video"]["size"][0]), int(cfg["video"]["size"][1])
        t_down, s_down = int(cfg["video"]["latent"]["t_down"]), int(cfg["video"]["latent"]["s_down"])
        shape_v = (B, Cv, max(1, T_in // t_down), H // s_down, W // s_down)
        z_init = torch.randn(*shape_v, device=device) if z_init is None else z_init.to(device).float()
        if tuple(z_init.shape) != shape_v:
            raise ValueError(f"z_init must have shape {shape_v}")
        target = "video"
    z = sample_latents(cfg=cfg, adapt_v=adapt_v, adapt_a=adapt_a, core=core, head=head, target=target,
                       z_prompt=z_prompt.to(device).float().contiguous(), z_init=z_init.contiguous(),
                       core_precision=core_precision, head_precision=head_precision)
    if target == "audio":
        wav_hat = aud_codec.decode(z).squeeze(0).squeeze(0).detach().cpu().numpy()
        return {"audio": wav_hat, "sr": sr}
    x_hat = vid_vae.decode(z).clamp(0, 1)
    x_hat = (x_hat[0].permute(1, 2, 3, 0).detach().cpu().numpy() * 255.0).astype(np.uint8)
    return {"video": x_hat, "fps": fps}
