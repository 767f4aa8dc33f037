"""ctypes binding of libavdiff_b200.so (C ABI declared in include/avdiff_b200.h).

This is the ONLY way the package reaches the GPU kernels.  There is no fallback: if the shared
library is missing or cannot be loaded, importing/using any operator raises ``AvdLibraryError``.
The structures below mirror include/avdiff_b200.h field for field.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional

import torch

PKG = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("AVD_LIB_PATH") or (PKG / "libavdiff_b200.so"))   # AVD_LIB_PATH: an experimental build (tools/exp_build.sh)

AVD_PREC_BF16 = 1
AVD_PREC_F16 = 2
AVD_F16_XSCALE = 1.0 / 256.0
AVD_PREC_BF16X3 = 3
AVD_PREC_FP32 = 6
AVD_TARGET_VIDEO = 0
AVD_TARGET_AUDIO = 1
AVD_MAX_LAYERS = 64
AVD_MAX_SHARED = 8
AVD_CHAIN_FLAG_BYTES = 2560
AVD_AUDIO_DECODE_STATIC_MAX_BYTES = 16 * 1024 * 1024   # include/avdiff_b200.h

PRECISIONS = {"bf16": AVD_PREC_BF16, "f16": AVD_PREC_F16, "bf16x3": AVD_PREC_BF16X3, "fp32": AVD_PREC_FP32}


class AvdLibraryError(RuntimeError):
    """libavdiff_b200.so is missing / not loadable (there is no CPU or PyTorch fallback)."""


class AvdError(RuntimeError):
    """A C-ABI call returned a negative status; the message is avd_last_error()."""


vp = C.c_void_p
fp = C.c_void_p      # const float* (device) -- passed as raw addresses
i64p = C.c_void_p


class CfgDdimArgs(C.Structure):
    _fields_ = [("target", C.c_int), ("B", C.c_int),
                ("C", C.c_int), ("T", C.c_int), ("H", C.c_int), ("W", C.c_int),
                ("tt", C.c_int), ("th", C.c_int), ("tw", C.c_int),
                ("L", C.c_int), ("length", C.c_int), ("stride", C.c_int), ("Nt", C.c_int),
                ("guidance", C.c_float), ("t_now", C.c_int), ("t_prev", C.c_int), ("T_train", C.c_int),
                ("eta", C.c_float), ("step", C.c_int), ("seed", C.c_uint64), ("clip0", C.c_int64)]


class GemmArgs(C.Structure):
    _fields_ = [("M", C.c_int), ("N", C.c_int), ("K", C.c_int), ("precision", C.c_int),
                ("A", vp), ("W", vp), ("bias", fp), ("ssq_in", fp), ("ssq_parts", C.c_int),
                ("rms_dim", C.c_int), ("rms_eps", C.c_float), ("resid", fp), ("act", C.c_int),
                ("out_f32", fp), ("out_bf16", vp), ("out_precision", C.c_int), ("ssq_out", fp),
                ("a_scale", C.c_float), ("out_scale", C.c_float)]


class BlockWeights(C.Structure):
    _fields_ = [("w_qkv", vp), ("b_qkv", fp), ("w_o", vp), ("b_o", fp),
                ("w_fc1", vp), ("b_fc1", fp), ("w_fc2", vp), ("b_fc2", fp),
                ("ln1_w", fp), ("ln1_b", fp), ("ln2_w", fp), ("ln2_b", fp)]


class CoreWeights(C.Structure):
    _fields_ = [("d_model", C.c_int), ("n_heads", C.c_int), ("n_layers", C.c_int), ("d_mlp", C.c_int),
                ("precision", C.c_int), ("rms_eps", C.c_float),
                ("blocks", BlockWeights * AVD_MAX_LAYERS), ("final_scale", fp),
                ("norm_kind", C.c_int), ("ln_eps", C.c_float), ("final_ln_w", fp), ("final_ln_b", fp)]


class CoreWorkspace(C.Structure):
    _fields_ = [("x", fp), ("xa", vp), ("ssq", fp), ("qkv", vp), ("att", vp), ("hmid", vp), ("key_mask", vp),
                ("f16_sat_count", vp), ("chain_flags", vp), ("profile_skip_kinds", C.c_int)]


class HeadWeights(C.Structure):
    _fields_ = [("d_in", C.c_int), ("hidden", C.c_int), ("d_out", C.c_int), ("n_shared", C.c_int),
                ("precision", C.c_int), ("ln_eps", C.c_float),
                ("w_in", vp), ("b_in", fp),
                ("w_s", vp * AVD_MAX_SHARED), ("b_s", fp * AVD_MAX_SHARED),
                ("ln_w", fp * AVD_MAX_SHARED), ("ln_b", fp * AVD_MAX_SHARED),
                ("w_out", vp), ("b_out", fp), ("activation", C.c_int)]


class SamplerArgs(C.Structure):
    _fields_ = [("target", C.c_int), ("B", C.c_int),
                ("Cv", C.c_int), ("T", C.c_int), ("H", C.c_int), ("W", C.c_int),
                ("tt", C.c_int), ("th", C.c_int), ("tw", C.c_int),
                ("Ca", C.c_int), ("Fa", C.c_int), ("chunk_len", C.c_int), ("chunk_stride", C.c_int),
                ("tstep_dim", C.c_int), ("steps", C.c_int), ("T_train", C.c_int), ("guidance", C.c_float),
                ("timesteps_host", i64p), ("temb_steps", fp), ("temb_zero", fp), ("alpha_bar", fp),
                ("wv_t", fp), ("bv", fp), ("wa_t", fp), ("ba", fp),
                ("core", CoreWeights), ("head", HeadWeights),
                ("workspace", vp), ("workspace_bytes", C.c_int64),
                ("eta", C.c_float), ("seed", C.c_uint64), ("clip_index0", C.c_int64),
                ("add_v", fp), ("add_a", fp), ("tadd_steps", fp), ("tadd_zero", fp),
                ("profile_skip_kinds", C.c_int), ("f16_sat_count", vp),
                ("wv_pack", vp), ("wa_pack", vp), ("adapt_precision", C.c_int)]


AVD_VAE_MAX_BLOCKS = 8


class VaeWeights(C.Structure):
    _fields_ = [("in_ch", C.c_int), ("lat_ch", C.c_int), ("base", C.c_int), ("t_down", C.c_int), ("s_down", C.c_int),
                ("enc_blocks", C.c_int), ("dec_blocks", C.c_int), ("out_tanh", C.c_int), ("gn_eps", C.c_float),
                ("enc_w", fp * AVD_VAE_MAX_BLOCKS), ("enc_b", fp * AVD_VAE_MAX_BLOCKS),
                ("enc_gn_w", fp * AVD_VAE_MAX_BLOCKS), ("enc_gn_b", fp * AVD_VAE_MAX_BLOCKS),
                ("to_lat_w", fp), ("to_lat_b", fp), ("from_lat_w", fp), ("from_lat_b", fp),
                ("dec_w", fp * AVD_VAE_MAX_BLOCKS), ("dec_b", fp * AVD_VAE_MAX_BLOCKS),
                ("dec_gn_w", fp * AVD_VAE_MAX_BLOCKS), ("dec_gn_b", fp * AVD_VAE_MAX_BLOCKS),
                ("to_img_w", fp), ("to_img_b", fp)]


# name -> (restype, argtypes); every symbol include/avdiff_b200.h declares
_I = C.c_int
_SIGNATURES = {
    "avd_last_error": (C.c_char_p, []),
    "avd_version": (_I, []),
    "avd_nseg": (_I, [_I]),
    "avd_device_check": (_I, []),
    "avd_struct_size": (_I, [_I]),
    "avd_prof_enable": (_I, [_I]),
    "avd_prof_reset": (_I, []),
    "avd_prof_num_kinds": (_I, []),
    "avd_prof_kind_name": (C.c_char_p, [_I]),
    "avd_prof_collect": (_I, [_I, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "avd_launch_count": (C.c_int64, []),
    "avd_debug_trace": (_I, [vp, C.c_int64]),
    "avd_debug_gemm_timeout": (_I, [C.POINTER(C.c_uint)]),
    "avd_check_errors": (_I, []),
    "avd_set_option": (_I, [C.c_char_p, _I]),
    "avd_get_option": (_I, [C.c_char_p, C.POINTER(_I)]),
    "avd_philox_normal": (_I, [fp, C.c_uint64, C.c_int64, _I, _I, C.c_int64, vp]),
    "avd_f16_range_scan": (_I, [vp, C.c_int64, vp, vp]),
    "avd_tube_patch": (_I, [fp, fp, _I, _I, _I, _I, _I, _I, _I, _I, vp]),
    "avd_tube_unpatch": (_I, [fp, fp, _I, _I, _I, _I, _I, _I, _I, _I, vp]),
    "avd_chunk1d": (_I, [fp, fp, _I, _I, _I, _I, _I, vp]),
    "avd_unchunk1d": (_I, [fp, fp, _I, _I, _I, _I, _I, _I, vp]),
    "avd_crossfade_audio": (_I, [fp, fp, fp, _I, _I, _I, vp]),
    "avd_overlap_add_win": (_I, [fp, fp, fp, _I, _I, _I, _I, C.c_float, vp]),
    "avd_crossfade_video": (_I, [vp, fp, vp, _I, _I, C.c_int64, _I, vp]),
    "avd_timestep_embedding": (_I, [fp, fp, _I, _I, C.c_float, vp]),
    "avd_ddim_step": (_I, [fp, fp, fp, fp, i64p, i64p, fp, _I, _I, C.c_int64, C.c_float, vp]),
    "avd_cfg_untok_ddim": (_I, [C.POINTER(CfgDdimArgs), fp, fp, fp, fp, fp, vp]),
    "avd_pack_weight": (_I, [fp, fp, vp, _I, _I, _I, vp]),
    "avd_transpose_f32": (_I, [fp, fp, _I, _I, vp]),
    "avd_rows_prepare": (_I, [fp, vp, fp, _I, _I, _I, vp]),
    "avd_rmsnorm_rows": (_I, [fp, fp, _I, _I, _I, _I, _I, C.c_float, fp, vp, _I, vp]),
    "avd_layernorm_act": (_I, [fp, fp, fp, _I, _I, C.c_float, _I, fp, vp, _I, vp]),
    "avd_gemm": (_I, [C.POINTER(GemmArgs), vp]),
    "avd_attention": (_I, [vp, _I, vp, _I, _I, _I, _I, _I, vp]),
    "avd_attention_masked": (_I, [vp, _I, vp, _I, _I, _I, _I, _I, vp, vp]),
    "avd_mmdt_blocks": (_I, [C.POINTER(CoreWeights), fp, vp, fp, C.POINTER(CoreWorkspace), _I, _I, vp]),
    "avd_head_forward": (_I, [C.POINTER(HeadWeights), vp, _I, vp, vp, fp, fp, vp]),
    "avd_sampler_workspace_bytes": (C.c_int64, [C.POINTER(SamplerArgs)]),
    "avd_sample_prepare": (_I, [C.POINTER(SamplerArgs), fp, vp]),
    "avd_sample_step": (_I, [C.POINTER(SamplerArgs), _I, fp, fp, vp]),
    "avd_sample_loop": (_I, [C.POINTER(SamplerArgs), fp, fp, vp]),
    "avd_audio_encode": (_I, [fp, fp, _I, _I, _I, _I, _I, _I, fp, fp, fp, fp, fp, fp, vp]),
    "avd_audio_decode_scratch_bytes": (C.c_int64, [_I, _I, _I, _I]),
    "avd_audio_decode": (_I, [fp, fp, fp, _I, _I, _I, _I, _I, _I, fp, fp, fp, fp, fp, fp, fp, fp, vp]),
    "avd_audio_decode_static": (_I, [fp, fp, _I, _I, _I, _I, _I, _I, fp, fp, fp, fp, fp, fp, fp, fp, vp]),
    "avd_vae_scratch_bytes": (C.c_int64, [_I, _I, _I, _I]),
    "avd_video_encode": (_I, [C.POINTER(VaeWeights), fp, fp, vp, _I, _I, _I, _I, vp]),
    "avd_video_decode": (_I, [C.POINTER(VaeWeights), fp, fp, vp, _I, _I, _I, _I, _I, _I, _I, vp]),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)
ABI_STRUCTS = (CfgDdimArgs, GemmArgs, BlockWeights, CoreWeights, CoreWorkspace, HeadWeights, SamplerArgs, VaeWeights)

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Load libavdiff_b200.so (built by ``python -m multimodal_diffusion_b200.build``).  No fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise AvdLibraryError(
            f"{LIB_PATH} is missing: build it with `python -m multimodal_diffusion_b200.build` "
            "(nvcc, sm_100a).  multimodal_diffusion_b200 has no CPU / PyTorch fallback.")
    try:
        lib = C.CDLL(str(LIB_PATH))
    except OSError as e:  # pragma: no cover - depends on the host
        raise AvdLibraryError(f"cannot load {LIB_PATH}: {e}") from e
    for name, (res, args) in _SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise AvdLibraryError(f"{LIB_PATH} does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    for i, st in enumerate(ABI_STRUCTS):          # the ctypes mirrors must match the compiled structs byte for byte
        if lib.avd_struct_size(i) != C.sizeof(st):
            raise AvdLibraryError(f"ABI mismatch: sizeof({st.__name__}) is {C.sizeof(st)} in the binding but "
                                  f"{lib.avd_struct_size(i)} in {LIB_PATH.name}; rebuild the library")
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise AvdError(load().avd_last_error().decode("utf-8", "replace"))


def stream_ptr(device=None) -> int:
    """The raw cudaStream_t of torch's current stream on `device` (default: the current device)."""
    return torch.cuda.current_stream(device).cuda_stream


def set_option(name: str, value: int) -> None:
    """Tuning switch of the library (include/avdiff_b200.h: avd_set_option); the AVD_* environment variables are read
    once at load, later changes go through here."""
    check(load().avd_set_option(name.encode(), int(value)))


def get_option(name: str) -> int:
    v = C.c_int(0)
    check(load().avd_get_option(name.encode(), C.byref(v)))
    return int(v.value)


class option:
    """Context manager: `with _lib.option("gemm_pair", 2): ...` sets a tuning switch and restores it."""

    def __init__(self, name: str, value: int):
        self.name, self.value = name, value

    def __enter__(self):
        self.prev = get_option(self.name)
        set_option(self.name, self.value)
        return self

    def __exit__(self, *exc):
        set_option(self.name, self.prev)


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def require_cuda_f32(t: torch.Tensor, name: str) -> torch.Tensor:
    """The C ABI takes raw device pointers to contiguous fp32 data; validate before any launch."""
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise AvdError(f"{name} must live on a CUDA device (multimodal_diffusion_b200 has no CPU path)")
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


def precision_code(p) -> int:
    if isinstance(p, str):
        if p not in PRECISIONS:
            raise ValueError(f"precision must be one of {sorted(PRECISIONS)}, got {p!r}")
        return PRECISIONS[p]
    if p in (AVD_PREC_BF16, AVD_PREC_F16, AVD_PREC_BF16X3, AVD_PREC_FP32):
        return int(p)
    raise ValueError(f"bad precision {p!r}")


def nseg(p) -> int:
    return {1: 1, 2: 1, 3: 2, 6: 3}[precision_code(p)]
