"""The fused DDIM + CFG sampling loop (K8): host-side owner of the buffers that avd_sample_* work on.

One ``FusedSampler`` = one direction (target modality), one clip-batch size B, one set of packed weights.
Per DDIM step the C ABI enqueues:
    build_x (tokenise target latent + adapter + concat timestep embedding, cond and null sequences at once)
    -> 8 x [QKV GEMM, joint attention, out-proj GEMM(+residual,+row ssq), fc1 GEMM(+GELU), fc2 GEMM(+residual,+row ssq)]
    -> final RMSNorm on the target rows -> head (4 GEMMs + 2 LayerNorm/GELU) -> fused CFG + un-tokenise + DDIM.
The cond / null CFG branches are co-batched as 2B sequences, the prompt side is built once
(F10), the non-target head is skipped (F9).  The whole S-step loop is captured into one CUDA graph.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib, packing
from .avdiff.utils import schedule_utils as su


@dataclass
class HotPathConfig:
    """The scalars sample_one_direction reads from the merged YAML (sample_clip.py:241-265); defaults = mvp.yaml
    merged with a2v.yaml / v2a.yaml, with the benchmark's 50 sampler steps."""
    d_model: int = 512
    n_layers: int = 8
    n_heads: int = 8
    mlp_ratio: float = 4.0
    tstep_dim: int = 256
    head_hidden: int = 512
    Cv: int = 8
    tube: Tuple[int, int, int] = (2, 4, 4)      # the sampler uses tube.h for both h and w (sample_clip.py:252,181)
    Ca: int = 8
    chunk: Tuple[int, int] = (4, 4)             # (length, stride)
    Fa: int = 150
    video_latent: Tuple[int, int, int] = (12, 16, 16)   # T', H', W' (48 frames / 4, 128 / 8)
    T_train: int = 1000
    schedule: str = "cosine"
    min_beta: float = 1e-4
    max_beta: float = 0.02
    sampler_steps: int = 50
    guidance: float = 3.5
    eta: float = 0.0

    @classmethod
    def from_yaml_dict(cls, cfg: Dict, target: str, sampler_steps: Optional[int] = None) -> "HotPathConfig":
        """Same keys as sample_clip.py:78-106,241-265 on the deep-merged config dict."""
        core = cfg["model"]["core"]
        dif = cfg["diffusion"][target]
        p = int(cfg["tokenizer"]["video"]["tube"]["h"])
        fps = int(cfg["video"]["fps"])
        T_in = int(round(cfg["data"]["clip_seconds"] * fps)) if "data" in cfg and "clip_seconds" in cfg["data"] else 48
        H, W = int(cfg["video"]["size"][0]), int(cfg["video"]["size"][1])
        t_down, s_down = int(cfg["video"]["latent"]["t_down"]), int(cfg["video"]["latent"]["s_down"])
        return cls(
            d_model=int(core["d_model"]), n_layers=int(core["n_layers"]), n_heads=int(core["n_heads"]),
            mlp_ratio=float(core.get("mlp_ratio", 4.0)),
            tstep_dim=int(cfg["embeddings"].get("timestep_dim", 256)),
            head_hidden=int(cfg["model"]["heads"]["video"]["hidden_dim"]),
            Cv=int(cfg["video"]["latent"]["channels"]),
            tube=(int(cfg["tokenizer"]["video"]["tube"]["t"]), p, p),
            Ca=int(cfg["audio"]["latent"]["channels"]),
            chunk=(int(cfg["tokenizer"]["audio"]["chunk"]["length"]), int(cfg["tokenizer"]["audio"]["chunk"]["stride"])),
            Fa=int(cfg["audio"]["latent"]["frames_per_clip"]),
            video_latent=(max(1, T_in // t_down), H // s_down, W // s_down),
            T_train=int(dif["steps"]), schedule=dif["schedule"], min_beta=float(dif["min_beta"]),
            max_beta=float(dif["max_beta"]),
            sampler_steps=int(sampler_steps if sampler_steps is not None else dif["sampler_steps"]),
            guidance=float(cfg["sampling"]["guidance_scale"].get(target, 3.0)),
            eta=float(cfg["sampling"].get("ddim_eta", 0.0)))

    # derived shapes
    @property
    def Nv(self) -> int:
        T, H, W = self.video_latent
        return (T // self.tube[0]) * (H // self.tube[1]) * (W // self.tube[2])

    @property
    def Na(self) -> int:
        return (self.Fa - self.chunk[0]) // self.chunk[1] + 1

    @property
    def Dv(self) -> int:
        return self.Cv * self.tube[0] * self.tube[1] * self.tube[2]

    @property
    def Da(self) -> int:
        return self.Ca * self.chunk[0]

    def flops_per_clip(self, target: str) -> float:
        """Algorithmic FLOPs of one clip's S-step sampling: 2 core forwards + 2 head passes over
        the target tokens + 1 target adapter per step."""
        d, L, N = self.d_model, self.n_layers, self.Nv + self.Na
        core = L * N * (24 * d * d + 4 * N * d) if self.mlp_ratio == 4.0 else \
            L * N * ((8 + 4 * self.mlp_ratio) * d * d + 4 * N * d)
        Nt, Dt = (self.Nv, self.Dv) if target == "video" else (self.Na, self.Da)
        hid = self.head_hidden
        head_tok = 2 * (d * hid + 2 * hid * hid + hid * Dt)
        adapt_tok = 2 * Dt * (d - self.tstep_dim)
        return self.sampler_steps * (2 * core + 2 * Nt * head_tok + Nt * adapt_tok)


def _state_dict_of(obj) -> Dict[str, torch.Tensor]:
    return obj if isinstance(obj, dict) else obj.state_dict()


# The ONE default precision of every sampling entry point (FusedSampler, sample_latents, sample_one_direction,
# sample_sharded, bench.py): fp16 operands / fp32 accumulation for the core and the head -- the timed configuration and
# the 16-bit mode that meets every tolerance of the hot-path spec.  Stand-alone module calls
# (MMDiT.forward, the head, the adapters) default to the fp32-parity mode instead; see avdiff/models/mmdt.py.
DEFAULT_CORE_PRECISION = "f16"
DEFAULT_HEAD_PRECISION = "f16"


class AvdRangeError(_lib.AvdError):
    """fp16 mode: a 16-bit operand saturated (or went non-finite) -- the result cannot be trusted; use bf16 / fp32."""


class FusedSampler:
    """sample_one_direction's loop (sample_clip.py:317-348 / 359-389) for a batch of B clips on one GPU.

    adapt_v / adapt_a / core / head: reference (or mirror) modules or their ``state_dict``s.
    target: "audio" (V2A: z_prompt = video latent [B,Cv,T',H',W'], z = audio latent [B,Ca,Fa]) or
            "video" (A2V: z_prompt = audio latent, z = video latent).
    activation: the head's activation when `head` is a bare state_dict (read from the module otherwise).
    seed / clip_index0: eta > 0 only -- Philox key and the global index of clip 0 of this batch (shard offset).
    add_v / add_a: optional per-token rows [Nv, d - tstep_dim] / [Na, d - tstep_dim] added after the adapters
            (positional + modality embeddings, adapters.py:41-127).
    time_embedding: "concat" (sample_clip.py:59-70) or "add" (trainer.py:45-49: adapters project to the full width d and
            the sinusoidal embedding of width d is ADDED; cfg.tstep_dim is then ignored).
    """

    def __init__(self, cfg: HotPathConfig, adapt_v, adapt_a, core, head, target: str, batch: int,
                 device="cuda", core_precision: str = DEFAULT_CORE_PRECISION,
                 head_precision: str = DEFAULT_HEAD_PRECISION, activation: Optional[str] = None, seed: int = 0,
                 clip_index0: int = 0, add_v: Optional[torch.Tensor] = None, add_a: Optional[torch.Tensor] = None,
                 time_embedding: str = "concat"):
        if target not in ("audio", "video"):
            raise ValueError("target must be 'audio' or 'video'")
        if time_embedding not in ("concat", "add"):
            raise ValueError("time_embedding must be 'concat' or 'add'")
        if cfg.eta < 0.0:
            raise ValueError("ddim_eta must be >= 0")
        lib = _lib.load()
        self.cfg, self.target, self.B = cfg, target, int(batch)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.AvdError("FusedSampler needs a CUDA device (no CPU path)")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        # hyper-parameters a state_dict does not carry come from the live modules
        hy_core, hy_head = packing.module_hyper(core), packing.module_hyper(head)
        for name, hy in (("core", hy_core), ("head", hy_head)):
            if hy.get("training") and hy.get("dropout"):
                raise NotImplementedError(f"{name} is in training mode with dropout > 0 (the reference would apply it); "
                                          "call .eval() -- the fused sampler is the eval-mode path (sample_clip.py:98,108)")
        if activation is None:
            activation = hy_head.get("activation")
            if activation is None and isinstance(head, dict):
                raise ValueError("head given as a bare state_dict: pass activation='gelu' | 'relu' | 'leaky_relu' "
                                 "(cfg.model.heads.video.activation)")
        tdim = cfg.tstep_dim if time_embedding == "concat" else 0
        with torch.cuda.device(self.device):
            _lib.check(lib.avd_device_check())
            sd_core = _state_dict_of(core)
            self.core = packing.PackedCore(sd_core, cfg.n_layers, cfg.n_heads, core_precision, self.device,
                                           rms_eps=float(hy_core.get("rms_eps", 1e-6)),
                                           ln_eps=float(hy_core.get("ln_eps", 1e-5)))
            # rmsnorm: final_norm.scale is folded into the head's input projection (the norm itself runs in that GEMM's
            # epilogue); layernorm: the final nn.LayerNorm is its own pass and the head sees normalised rows
            self.head = packing.PackedHead(_state_dict_of(head), target, head_precision, self.device,
                                           ln_eps=float(hy_head.get("ln_eps", 1e-5)),
                                           in_col_scale=None if self.core.layernorm else sd_core["final_norm.scale"],
                                           activation=activation or "gelu")
            # the adapter's tensor-core route keeps ~fp32 products: 6 split-bf16 passes next to an fp32 core, 3 otherwise
            self.adapt_precision = "fp32" if _lib.precision_code(core_precision) == _lib.AVD_PREC_FP32 else "bf16x3"
            self.av = packing.PackedAdapter(_state_dict_of(adapt_v), self.device, tc_precision=self.adapt_precision)
            self.aa = packing.PackedAdapter(_state_dict_of(adapt_a), self.device, tc_precision=self.adapt_precision)
        d_adapt = cfg.d_model - tdim
        if self.av.d_in != cfg.Dv or self.aa.d_in != cfg.Da or self.av.d_out != d_adapt or self.aa.d_out != d_adapt:
            raise ValueError(f"adapter shapes do not match the config (expected {cfg.Dv}/{cfg.Da} -> {d_adapt}; a trainer "
                             "checkpoint's full-width adapters need time_embedding='add')")

        # host tables (reference functions' fp32 results)
        alpha_bar = su.alphas_cumprod_from_betas(
            su.make_beta_schedule(cfg.T_train, cfg.schedule, cfg.min_beta, cfg.max_beta))[1]
        self.timesteps = su.make_sampling_schedule(cfg.T_train, cfg.sampler_steps)        # int64 [S+1], host
        self._ts_host = np.ascontiguousarray(self.timesteps.numpy().astype(np.int64))
        self.alpha_bar = alpha_bar.to(self.device)
        emb_dim = cfg.tstep_dim if time_embedding == "concat" else cfg.d_model
        self.temb_steps = su.timestep_embedding_host(self.timesteps[:-1], emb_dim).to(self.device).contiguous()
        self.temb_zero = su.timestep_embedding_host(torch.zeros(1, dtype=torch.long), emb_dim)[0] \
            .to(self.device).contiguous()
        self.add_v = None if add_v is None else packing._dev_f32(add_v, self.device)
        self.add_a = None if add_a is None else packing._dev_f32(add_a, self.device)
        if self.add_v is not None and tuple(self.add_v.shape) != (cfg.Nv, d_adapt):
            raise ValueError(f"add_v must have shape {(cfg.Nv, d_adapt)}")
        if self.add_a is not None and tuple(self.add_a.shape) != (cfg.Na, d_adapt):
            raise ValueError(f"add_a must have shape {(cfg.Na, d_adapt)}")
        self.sat_count = torch.zeros(1, device=self.device, dtype=torch.int32)     # fp16 range-check counter

        s = _lib.SamplerArgs()
        s.target = _lib.AVD_TARGET_AUDIO if target == "audio" else _lib.AVD_TARGET_VIDEO
        s.B = self.B
        s.Cv = cfg.Cv
        s.T, s.H, s.W = cfg.video_latent
        s.tt, s.th, s.tw = cfg.tube
        s.Ca, s.Fa, s.chunk_len, s.chunk_stride = cfg.Ca, cfg.Fa, cfg.chunk[0], cfg.chunk[1]
        s.tstep_dim, s.steps, s.T_train, s.guidance = tdim, cfg.sampler_steps, cfg.T_train, cfg.guidance
        s.timesteps_host = self._ts_host.ctypes.data
        if time_embedding == "concat":
            s.temb_steps, s.temb_zero = self.temb_steps.data_ptr(), self.temb_zero.data_ptr()
        else:
            s.tadd_steps, s.tadd_zero = self.temb_steps.data_ptr(), self.temb_zero.data_ptr()
        s.alpha_bar = self.alpha_bar.data_ptr()
        s.wv_t, s.bv = self.av.wt.data_ptr(), self.av.bias.data_ptr()
        s.wa_t, s.ba = self.aa.wt.data_ptr(), self.aa.bias.data_ptr()
        s.add_v, s.add_a = _lib.ptr(self.add_v), _lib.ptr(self.add_a)
        s.wv_pack, s.wa_pack = _lib.ptr(self.av.pack), _lib.ptr(self.aa.pack)
        s.adapt_precision = _lib.precision_code(self.adapt_precision)
        s.core, s.head = self.core.struct, self.head.struct
        s.eta, s.seed, s.clip_index0 = float(cfg.eta), int(seed) & (2 ** 64 - 1), int(clip_index0)
        s.workspace, s.workspace_bytes = None, 0
        nbytes = lib.avd_sampler_workspace_bytes(C.byref(s))
        if nbytes < 0:
            raise _lib.AvdError(lib.avd_last_error().decode())
        self.workspace = torch.empty(int(nbytes), device=self.device, dtype=torch.uint8)
        s.workspace, s.workspace_bytes = self.workspace.data_ptr(), int(nbytes)
        self.args = s
        self.Nt, self.Dt = (cfg.Na, cfg.Da) if target == "audio" else (cfg.Nv, cfg.Dv)
        self.prompt_shape = (self.B, cfg.Cv, *cfg.video_latent) if target == "audio" else (self.B, cfg.Ca, cfg.Fa)
        self.latent_shape = (self.B, cfg.Ca, cfg.Fa) if target == "audio" else (self.B, cfg.Cv, *cfg.video_latent)
        self.is_f16 = _lib.AVD_PREC_F16 in (_lib.precision_code(core_precision), _lib.precision_code(head_precision))
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._g_prompt: Optional[torch.Tensor] = None
        self._g_z: Optional[torch.Tensor] = None

    # ---- per-call API (eager launches on the current stream of self.device) -------------------------
    def _chk(self, t: torch.Tensor, shape, name: str) -> torch.Tensor:
        t = _lib.require_cuda_f32(t, name)
        if t.device != self.device:
            raise ValueError(f"{name} lives on {t.device}, the sampler on {self.device}")
        if tuple(t.shape) != tuple(shape):
            raise ValueError(f"{name} must have shape {tuple(shape)}, got {tuple(t.shape)}")
        return t

    def _stream(self) -> int:
        return _lib.stream_ptr(self.device)

    def set_clip_index0(self, clip_index0: int) -> None:
        """eta > 0: global index of clip 0 of the NEXT call (the Philox counter word).  Not valid after capture()."""
        if self._graph is not None and int(clip_index0) != int(self.args.clip_index0):
            raise RuntimeError("clip_index0 is baked into the captured graph; build another FusedSampler")
        self.args.clip_index0 = int(clip_index0)

    def prepare(self, z_prompt: torch.Tensor) -> None:
        zp = self._chk(z_prompt, self.prompt_shape, "z_prompt")
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().avd_sample_prepare(C.byref(self.args), zp.data_ptr(), self._stream()))

    def step(self, i: int, z: torch.Tensor, eps_tap: Optional[torch.Tensor] = None) -> None:
        """One DDIM+CFG step, z updated IN PLACE.  eps_tap [2B,Nt,Dt] receives (eps_cond ; eps_null)."""
        z = self._chk(z, self.latent_shape, "z")
        if eps_tap is not None:
            eps_tap = self._chk(eps_tap, (2 * self.B, self.Nt, self.Dt), "eps_tap")
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().avd_sample_step(C.byref(self.args), int(i), z.data_ptr(), _lib.ptr(eps_tap),
                                                   self._stream()))

    def sample_eager(self, z_prompt: torch.Tensor, z_init: torch.Tensor) -> torch.Tensor:
        zp = self._chk(z_prompt, self.prompt_shape, "z_prompt")
        z = self._chk(z_init, self.latent_shape, "z_init").clone()
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().avd_sample_loop(C.byref(self.args), zp.data_ptr(), z.data_ptr(), self._stream()))
        return z

    def sample_checked(self, z_prompt: torch.Tensor, z_init: torch.Tensor) -> torch.Tensor:
        """sample_eager with the fp16 range check on: every 16-bit operand buffer is scanned after each block and the
        head; raises AvdRangeError if any value saturated to +-65504 or went non-finite (the fixed 2^-8 residual-stream
        scale covers |x| < 1.6e7).  Also runs the library's asynchronous-error check.  Synchronises."""
        self.sat_count.zero_()
        self.args.f16_sat_count = self.sat_count.data_ptr()
        try:
            z = self.sample_eager(z_prompt, z_init)
            torch.cuda.synchronize(self.device)
        finally:
            self.args.f16_sat_count = None
        self.check_errors()
        n = int(self.sat_count.item())
        if n:
            raise AvdRangeError(f"{n} fp16 operand values saturated / non-finite during sampling: inputs outside the "
                                "range the fp16 mode covers; use core_precision='bf16' or 'fp32'")
        if not bool(torch.isfinite(z).all()):
            raise AvdRangeError("non-finite latent")
        return z

    def check_errors(self) -> None:
        """Hard-error check (barrier-timeout records of the persistent kernels, pending CUDA errors); call after a
        synchronise."""
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().avd_check_errors())

    # ---- CUDA-graph path -------------------------------------------------------------------------
    def capture(self) -> None:
        """Capture prepare + S steps into one CUDA graph over static prompt / latent buffers."""
        if self._graph is not None:
            return
        lib = _lib.load()
        with torch.cuda.device(self.device):
            self._g_prompt = torch.zeros(self.prompt_shape, device=self.device, dtype=torch.float32)
            self._g_z = torch.zeros(self.latent_shape, device=self.device, dtype=torch.float32)
            # warm-up launch outside capture (sets kernel attributes, builds the TMA descriptor cache)
            side = torch.cuda.Stream(device=self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):
                _lib.check(lib.avd_sample_prepare(C.byref(self.args), self._g_prompt.data_ptr(), self._stream()))
                _lib.check(lib.avd_sample_step(C.byref(self.args), 0, self._g_z.data_ptr(), None, self._stream()))
            torch.cuda.current_stream(self.device).wait_stream(side)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                _lib.check(lib.avd_sample_loop(C.byref(self.args), self._g_prompt.data_ptr(), self._g_z.data_ptr(),
                                               self._stream()))
        self._graph = g

    def sample(self, z_prompt: torch.Tensor, z_init: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """z_final for a batch (graph replay).  Inputs may be CUDA tensors or pinned host tensors."""
        self.capture()
        with torch.cuda.device(self.device):
            self._g_prompt.copy_(z_prompt.reshape(self.prompt_shape), non_blocking=True)
            self._g_z.copy_(z_init.reshape(self.latent_shape), non_blocking=True)
            self._graph.replay()
            if out is None:
                return self._g_z.clone()
            out.copy_(self._g_z, non_blocking=True)
            return out

    def kernels_per_sample(self) -> int:
        """Kernel launches of one S-step sampling call (the claim behind bench.py's gpu_launches)."""
        L, ns = self.cfg.n_layers, self.head.struct.n_shared
        per_step = 1 + (7 if self.core.layernorm else 5) * L + (2 + 2 * ns) + 1 + (1 if self.core.layernorm else 0)
        return 1 + self.cfg.sampler_steps * per_step     # + prompt-side build_x (memsets are not kernels of ours)


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block partition of clip indices over ranks: rank r owns [lo, hi)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def clip_noise(shape_per_clip, clip_index: int, base_seed: int) -> torch.Tensor:
    """Per-clip seeded N(0,1) latent (host generator) so results are invariant to how clips are sharded."""
    g = torch.Generator().manual_seed(int(base_seed) + int(clip_index))
    return torch.randn(*shape_per_clip, generator=g, dtype=torch.float32)


def sample_sharded(sampler_factory, z_prompt_all: torch.Tensor, z_init_all: torch.Tensor, *, rank: Optional[int] = None,
                   world: Optional[int] = None, group=None, gather: bool = True,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """The clip-sharded sampler (north_star item 4; the per-clip / per-window loops of stream_infer.py:182-188,207-213
    as ONE call): n clips are split into contiguous blocks over the ranks of `group` (`shard_range`), every rank runs
    its block through a FusedSampler, and the final latents are all-gathered (NCCL over NVLink; gloo for CPU tests of
    the host logic).  No collective on the data path besides that one gather.

    sampler_factory(batch, clip_index0) -> object with `.sample(z_prompt, z_init) -> z` and `.latent_shape`
        (normally `lambda b, i0: FusedSampler(cfg, ..., batch=b, clip_index0=i0)`); called once per distinct block size,
        cached on the function object by the caller if it wants to reuse graphs.
    z_prompt_all / z_init_all: [n, ...] host (or device) tensors holding ALL clips on every rank; each rank reads only
        its block.  Returns the [n, ...] final latents on every rank when gather=True, else this rank's block.
    Clips that do not divide evenly give the first ranks one more clip (blocks padded to the largest for the gather)."""
    import torch.distributed as dist
    if world is None:
        world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
    n = int(z_init_all.shape[0])
    lo, hi = shard_range(n, rank, world)
    nb = hi - lo
    max_nb = shard_range(n, 0, world)[1]
    z_local = None
    if nb > 0:
        fs = sampler_factory(nb, lo)
        dev = getattr(fs, "device", z_init_all.device)
        z_local = fs.sample(z_prompt_all[lo:hi].to(dev, non_blocking=True), z_init_all[lo:hi].to(dev, non_blocking=True))
    if not gather or world == 1:
        if out is not None and z_local is not None:
            out.copy_(z_local)
            return out
        return z_local
    per_clip = tuple(z_init_all.shape[1:])
    dev = z_local.device if z_local is not None else z_init_all.device
    pad = torch.zeros((max_nb,) + per_clip, device=dev, dtype=torch.float32)
    if nb > 0:
        pad[:nb].copy_(z_local)
    gathered = torch.empty((world, max_nb) + per_clip, device=dev, dtype=torch.float32)
    dist.all_gather_into_tensor(gathered.view(-1), pad.view(-1), group=group)
    if n == world * max_nb:                      # even split: the gathered buffer IS the result
        res = gathered.view((n,) + per_clip)
    else:
        res = torch.cat([gathered[r, :shard_range(n, r, world)[1] - shard_range(n, r, world)[0]] for r in range(world)])
    if out is not None:
        out.copy_(res)
        return out
    return res
