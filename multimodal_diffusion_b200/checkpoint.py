"""Checkpoint / weights interop.

The reference has two checkpoint dialects that do not meet:

* ``AVTrainer.save_checkpoint`` (avdiff/models/train/trainer.py:407-423) writes
  ``{"step", "core", "head", "adapt_v", "adapt_a", "vid_vae", "aud_codec", "opt", "ema"?}`` -- plain
  ``state_dict``s (DDP-wrapped modules are unwrapped on save, the EMA shadow of the core is NOT: its keys carry a
  ``module.`` prefix when the trainer ran under DDP, trainer.py:60-61,165).  The trainer's adapters project to the FULL
  width d and its timestep embedding of width d is ADDED (trainer.py:38-49,195-202).
* ``sample_clip.load_checkpoint_maybe`` (avdiff/models/infer/sample_clip.py:111-132) looks for
  ``{name}_state_dict`` keys (``core_state_dict`` ...) and the sampler's adapters project to d - tstep_dim with a
  CONCATENATED embedding (sample_clip.py:59-70,91-92).

``load_checkpoint`` reads either dialect (a path or an already loaded dict) and returns the four ``state_dict``s the
hot path needs plus which time-embedding variant the adapters imply; ``sampler_from_checkpoint`` builds a
``FusedSampler`` from it (``time_embedding="add"`` for trainer checkpoints: the fused input builder then adds the
width-d sinusoidal row instead of concatenating).  No module is instantiated: the packers work on raw tensors.
"""
from __future__ import annotations

from dataclasses import dataclass
from pathlib import Path
from typing import Dict, Optional, Union

import torch

PARTS = ("core", "head", "adapt_v", "adapt_a")


@dataclass
class HotPathWeights:
    core: Dict[str, torch.Tensor]
    head: Dict[str, torch.Tensor]
    adapt_v: Dict[str, torch.Tensor]
    adapt_a: Dict[str, torch.Tensor]
    time_embedding: str            # "concat" (sampler dialect) or "add" (trainer dialect)
    dialect: str                   # "trainer" | "sampler"
    used_ema: bool
    step: Optional[int] = None
    vid_vae: Optional[Dict[str, torch.Tensor]] = None
    aud_codec: Optional[Dict[str, torch.Tensor]] = None

    def as_dict(self) -> Dict[str, Dict[str, torch.Tensor]]:
        return {k: getattr(self, k) for k in PARTS}


def _strip_module_prefix(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """DistributedDataParallel prefixes every key with ``module.`` (the EMA shadow is taken from the wrapped core)."""
    if sd and all(k.startswith("module.") for k in sd):
        return {k[len("module."):]: v for k, v in sd.items()}
    return dict(sd)


def _d_model_of(core_sd: Dict[str, torch.Tensor]) -> int:
    for k in ("final_norm.scale", "final_norm.weight"):
        if k in core_sd:
            return int(core_sd[k].numel())
    raise ValueError("core state_dict has neither final_norm.scale (rmsnorm) nor final_norm.weight (layernorm)")


def load_checkpoint(src: Union[str, Path, Dict], use_ema: bool = True, map_location="cpu") -> HotPathWeights:
    """Read a trainer checkpoint (trainer.py:407-423) or a sampler-dialect one (sample_clip.py:111-132).

    use_ema: replace the core's weights by the EMA shadow when the checkpoint has one (the shadow tracks the core only,
    trainer.py:165; entries missing from it -- none in practice -- keep the raw weights)."""
    state = src if isinstance(src, dict) else torch.load(str(src), map_location=map_location, weights_only=False)
    if not isinstance(state, dict):
        raise ValueError("checkpoint is not a dict")
    if all(f"{p}_state_dict" in state for p in PARTS):
        dialect = "sampler"
        parts = {p: _strip_module_prefix(state[f"{p}_state_dict"]) for p in PARTS}
    elif all(p in state for p in PARTS):
        dialect = "trainer"
        parts = {p: _strip_module_prefix(state[p]) for p in PARTS}
    else:
        have = sorted(k for k in state if isinstance(k, str))
        raise ValueError(f"unrecognised checkpoint layout (keys: {have}); expected {PARTS} (trainer.py:408-416) or "
                         f"'<name>_state_dict' entries (sample_clip.py:122-126)")
    used_ema = False
    if use_ema and isinstance(state.get("ema"), dict) and state["ema"]:
        shadow = _strip_module_prefix(state["ema"])
        core = dict(parts["core"])
        for k, v in shadow.items():
            if k in core and core[k].shape == v.shape:
                core[k] = v
            elif k not in core:
                raise ValueError(f"EMA shadow entry {k!r} does not exist in the core state_dict")
        parts["core"] = core
        used_ema = True
    d = _d_model_of(parts["core"])
    w_v, w_a = parts["adapt_v"]["proj.weight"], parts["adapt_a"]["proj.weight"]
    if w_v.shape[0] != w_a.shape[0]:
        raise ValueError("video and audio adapters project to different widths")
    if int(w_v.shape[0]) == d:
        time_embedding = "add"          # trainer.py:45-49: full-width adapters, additive embedding
    elif 0 < int(w_v.shape[0]) < d:
        time_embedding = "concat"       # sample_clip.py:59-70: d - tstep_dim adapters, concatenated embedding
    else:
        raise ValueError(f"adapter width {int(w_v.shape[0])} does not fit d_model {d}")
    parts = {p: {k: v.detach().float() for k, v in sd.items()} for p, sd in parts.items()}
    return HotPathWeights(time_embedding=time_embedding, dialect=dialect, used_ema=used_ema,
                          step=int(state["step"]) if "step" in state else None,
                          vid_vae=state.get("vid_vae") or state.get("vid_vae_state_dict"),
                          aud_codec=state.get("aud_codec") or state.get("aud_codec_state_dict"), **parts)


def to_sampler_dialect(w: HotPathWeights) -> Dict[str, Dict[str, torch.Tensor]]:
    """The dict sample_clip.load_checkpoint_maybe expects (``<name>_state_dict`` keys), e.g. to re-save a trainer
    checkpoint for the reference's sampler.  (The reference sampler still cannot USE full-width adapters: F12.)"""
    return {f"{p}_state_dict": getattr(w, p) for p in PARTS}


def sampler_from_checkpoint(cfg, src: Union[str, Path, Dict, HotPathWeights], target: str, batch: int, *,
                            use_ema: bool = True, activation: str = "gelu", device="cuda", **kw):
    """FusedSampler over the weights of a checkpoint.  cfg: HotPathConfig; activation: the head's activation
    (cfg.model.heads.video.activation; not stored in a checkpoint)."""
    from .sampler import FusedSampler
    w = src if isinstance(src, HotPathWeights) else load_checkpoint(src, use_ema=use_ema)
    return FusedSampler(cfg, w.adapt_v, w.adapt_a, w.core, w.head, target, batch, device=device, activation=activation,
                        time_embedding=w.time_embedding, **kw)
