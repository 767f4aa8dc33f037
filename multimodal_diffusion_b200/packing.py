"""Weight import: reference ``state_dict`` tensors -> packed device operands of libavdiff_b200.so.

Layout in HBM: every nn.Linear weight [N,K] fp32 becomes ``nseg(precision)`` bf16 matrices
[nseg][N][K] (K contiguous = "K-major", the layout the TMA descriptors of the tcgen05 GEMM want); RMSNorm
``scale`` vectors are folded into the columns of the Linear that consumes the normalised activations; biases
stay fp32.  All buffers are torch CUDA tensors owned by the returned object (the library keeps no state).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch

from . import _lib


def _dev_f32(t: torch.Tensor, device) -> torch.Tensor:
    return t.detach().to(device=device, dtype=torch.float32).contiguous()


def pack_linear_weight(W: torch.Tensor, precision: int, col_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[N,K] fp32 (CUDA) -> bf16 [nseg,N,K] via avd_pack_weight (optionally W[:,k] *= col_scale[k])."""
    W = _lib.require_cuda_f32(W, "W")
    N, K = W.shape
    S = _lib.nseg(precision)
    out = torch.empty(S, N, K, device=W.device, dtype=torch.bfloat16)
    cs = None if col_scale is None else _lib.require_cuda_f32(col_scale, "col_scale")
    _lib.check(_lib.load().avd_pack_weight(W.data_ptr(), _lib.ptr(cs), out.data_ptr(), N, K, precision,
                                           _lib.stream_ptr()))
    return out


def transpose_f32(W: torch.Tensor) -> torch.Tensor:
    W = _lib.require_cuda_f32(W, "W")
    N, K = W.shape
    out = torch.empty(K, N, device=W.device, dtype=torch.float32)
    _lib.check(_lib.load().avd_transpose_f32(W.data_ptr(), out.data_ptr(), N, K, _lib.stream_ptr()))
    return out


ACTIVATIONS = {"gelu": 1, "leaky_relu": 2, "relu": 3}     # noise_heads._make_activation's names -> kernel code


def activation_code(name: Optional[str]) -> int:
    key = (name or "gelu").lower()
    if key not in ACTIVATIONS:
        raise ValueError(f"Unsupported activation: {name}")          # noise_heads.py:31
    return ACTIVATIONS[key]


def module_hyper(module) -> Dict[str, object]:
    """Hyper-parameters the packers cannot see in a state_dict, read from a live (reference or mirror) module: norm eps
    values, head activation, training flag.  Empty for a bare state_dict."""
    out: Dict[str, object] = {}
    if isinstance(module, dict) or module is None:
        return out
    out["training"] = bool(getattr(module, "training", False))
    for m in module.modules():
        cls = type(m).__name__
        if cls == "RMSNorm" and "rms_eps" not in out:
            out["rms_eps"] = float(getattr(m, "eps", 1e-6))
        elif cls == "LayerNorm" and "ln_eps" not in out:
            out["ln_eps"] = float(m.eps)
        elif cls in ("GELU", "LeakyReLU", "ReLU") and "activation" not in out:
            out["activation"] = {"GELU": "gelu", "LeakyReLU": "leaky_relu", "ReLU": "relu"}[cls]
        elif cls == "Dropout" and float(m.p) > 0.0:
            out["dropout"] = True
    return out


class PackedCore:
    """MMDiT weights (mmdt.py state_dict keys) packed for avd_mmdt_blocks."""

    def __init__(self, sd: Dict[str, torch.Tensor], n_layers: int, n_heads: int, precision, device,
                 rms_eps: float = 1e-6, ln_eps: float = 1e-5):
        self.precision = _lib.precision_code(precision)
        self.n_layers, self.n_heads = int(n_layers), int(n_heads)
        self.layernorm = "final_norm.scale" not in sd          # build_norm (mmdt.py:45-46): nn.LayerNorm has weight/bias
        if self.layernorm and "final_norm.weight" not in sd:
            raise ValueError("state_dict has neither final_norm.scale (rmsnorm) nor final_norm.weight (layernorm)")
        if self.n_layers > _lib.AVD_MAX_LAYERS:
            raise ValueError(f"at most {_lib.AVD_MAX_LAYERS} layers")
        self.keep: List[torch.Tensor] = []
        w = _lib.CoreWeights()
        d = int(sd["final_norm.weight" if self.layernorm else "final_norm.scale"].numel())
        self.d_model = d
        self.d_mlp = int(sd["blocks.0.mlp.fc1.weight"].shape[0])
        w.d_model, w.n_heads, w.n_layers, w.d_mlp = d, self.n_heads, self.n_layers, self.d_mlp
        w.precision, w.rms_eps = self.precision, float(rms_eps)
        for i in range(self.n_layers):
            p = f"blocks.{i}."
            s1 = None if self.layernorm else _dev_f32(sd[p + "norm1.scale"], device)
            s2 = None if self.layernorm else _dev_f32(sd[p + "norm2.scale"], device)
            wq = pack_linear_weight(_dev_f32(sd[p + "attn.mha.in_proj_weight"], device), self.precision, s1)
            wo = pack_linear_weight(_dev_f32(sd[p + "attn.mha.out_proj.weight"], device), self.precision)
            w1 = pack_linear_weight(_dev_f32(sd[p + "mlp.fc1.weight"], device), self.precision, s2)
            w2 = pack_linear_weight(_dev_f32(sd[p + "mlp.fc2.weight"], device), self.precision)
            bq = _dev_f32(sd[p + "attn.mha.in_proj_bias"], device)
            bo = _dev_f32(sd[p + "attn.mha.out_proj.bias"], device)
            b1 = _dev_f32(sd[p + "mlp.fc1.bias"], device)
            b2 = _dev_f32(sd[p + "mlp.fc2.bias"], device)
            self.keep += [wq, wo, w1, w2, bq, bo, b1, b2]
            bw = w.blocks[i]
            bw.w_qkv, bw.b_qkv = wq.data_ptr(), bq.data_ptr()
            bw.w_o, bw.b_o = wo.data_ptr(), bo.data_ptr()
            bw.w_fc1, bw.b_fc1 = w1.data_ptr(), b1.data_ptr()
            bw.w_fc2, bw.b_fc2 = w2.data_ptr(), b2.data_ptr()
            if self.layernorm:
                lnp = [_dev_f32(sd[p + k], device) for k in ("norm1.weight", "norm1.bias", "norm2.weight", "norm2.bias")]
                self.keep += lnp
                bw.ln1_w, bw.ln1_b, bw.ln2_w, bw.ln2_b = (t.data_ptr() for t in lnp)
        w.norm_kind, w.ln_eps = (1 if self.layernorm else 0), float(ln_eps)
        if self.layernorm:
            self.final_ln_w = _dev_f32(sd["final_norm.weight"], device)
            self.final_ln_b = _dev_f32(sd["final_norm.bias"], device)
            w.final_ln_w, w.final_ln_b = self.final_ln_w.data_ptr(), self.final_ln_b.data_ptr()
            self.final_scale = torch.ones(d, device=device, dtype=torch.float32)
        else:
            self.final_scale = _dev_f32(sd["final_norm.scale"], device)
        w.final_scale = self.final_scale.data_ptr()
        self.struct = w
        self.device = device

    def bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in self.keep) + self.final_scale.numel() * 4


class PackedHead:
    """One modality's path of MultiModalNoiseHead (noise_heads.py state_dict keys) for avd_head_forward:
    input_proj[m] -> shared[i] ... -> spec[m][j] (or shared_specific_trunk[j]) ... -> out_proj[m]."""

    def __init__(self, sd: Dict[str, torch.Tensor], modality: str, precision, device, ln_eps: float = 1e-5,
                 in_col_scale: Optional[torch.Tensor] = None, activation: Optional[str] = "gelu"):
        """in_col_scale: RMSNorm ``scale`` of the layer feeding the head (MMDiT.final_norm); when given it is folded
        into the columns of input_proj so the fused sampler can apply the norm as a row scale in the GEMM epilogue.
        activation: the head's `activation` argument ("gelu" / "relu" / "leaky_relu"); it is not part of a state_dict."""
        self.precision = _lib.precision_code(precision)
        self.modality = modality
        trunk: List[str] = []                                  # state_dict prefixes of the Linear+LN blocks, in order
        i = 0
        while f"shared.{i}.0.weight" in sd:
            trunk.append(f"shared.{i}")
            i += 1
        for prefix in (f"spec.{modality}", "shared_specific_trunk"):     # noise_heads.py:149-172, 213-218
            j = 0
            while f"{prefix}.{j}.0.weight" in sd:
                trunk.append(f"{prefix}.{j}")
                j += 1
        if len(trunk) > _lib.AVD_MAX_SHARED:
            raise ValueError(f"at most {_lib.AVD_MAX_SHARED} trunk layers (shared + modality-specific)")
        self.keep: List[torch.Tensor] = []
        w = _lib.HeadWeights()
        win = _dev_f32(sd[f"input_proj.{modality}.weight"], device)
        wout = _dev_f32(sd[f"out_proj.{modality}.weight"], device)
        w.hidden, w.d_in = int(win.shape[0]), int(win.shape[1])
        w.d_out, w.n_shared = int(wout.shape[0]), len(trunk)
        w.precision, w.ln_eps = self.precision, float(ln_eps)
        w.activation = activation_code(activation)
        cs = None if in_col_scale is None else _dev_f32(in_col_scale, device)
        pin, pout = pack_linear_weight(win, self.precision, cs), pack_linear_weight(wout, self.precision)
        bin_, bout = _dev_f32(sd[f"input_proj.{modality}.bias"], device), _dev_f32(sd[f"out_proj.{modality}.bias"], device)
        self.keep += [pin, pout, bin_, bout]
        w.w_in, w.b_in, w.w_out, w.b_out = pin.data_ptr(), bin_.data_ptr(), pout.data_ptr(), bout.data_ptr()
        for i, prefix in enumerate(trunk):
            ws = pack_linear_weight(_dev_f32(sd[f"{prefix}.0.weight"], device), self.precision)
            bs = _dev_f32(sd[f"{prefix}.0.bias"], device)
            lw = _dev_f32(sd[f"{prefix}.1.weight"], device)
            lb = _dev_f32(sd[f"{prefix}.1.bias"], device)
            self.keep += [ws, bs, lw, lb]
            w.w_s[i], w.b_s[i], w.ln_w[i], w.ln_b[i] = ws.data_ptr(), bs.data_ptr(), lw.data_ptr(), lb.data_ptr()
        self.d_in, self.hidden, self.d_out = w.d_in, w.hidden, w.d_out
        self.struct = w
        self.device = device


class PackedAdapter:
    """LinearAdapter (proj.weight [d_out,d_in], proj.bias): transposed fp32 for the FMA input builder, and (when the
    shape allows a tcgen05 GEMM: d_in % 64 == 0, d_out % 32 == 0) split-bf16 segments for its tensor-core route."""

    def __init__(self, sd: Dict[str, torch.Tensor], device, tc_precision=None):
        W = _dev_f32(sd["proj.weight"], device)
        self.wt = transpose_f32(W)            # [d_in, d_out]
        self.bias = _dev_f32(sd["proj.bias"], device)
        self.d_out, self.d_in = int(W.shape[0]), int(W.shape[1])
        self.pack = None
        if tc_precision is not None and self.d_in % 64 == 0 and self.d_out % 32 == 0:
            self.pack = pack_linear_weight(W, _lib.precision_code(tc_precision))


# ---------------------------------------------------------------------------------------------------
# small launch helpers shared by the module mirrors
# ---------------------------------------------------------------------------------------------------

def rows_prepare(x2d: torch.Tensor, precision: int):
    """x [M,d] fp32 -> (bf16 segments [S,M,d], ssq [M,d/64]) -- the A operand + RMSNorm row statistics."""
    M, d = x2d.shape
    if d % 64:
        raise ValueError("feature width must be a multiple of 64")
    S = _lib.nseg(precision)
    xa = torch.empty(S, M, d, device=x2d.device, dtype=torch.bfloat16)
    ssq = torch.empty(M, d // 64, device=x2d.device, dtype=torch.float32)
    if M > 0:
        _lib.check(_lib.load().avd_rows_prepare(x2d.data_ptr(), xa.data_ptr(), ssq.data_ptr(), M, d, precision,
                                                _lib.stream_ptr()))
    return xa, ssq


def gemm(A: torch.Tensor, W: torch.Tensor, M: int, N: int, K: int, precision: int, *, bias=None, ssq_in=None,
         rms_dim: int = 0, rms_eps: float = 0.0, resid=None, act: int = 0, out_f32=None, out_bf16=None,
         out_precision: int = 0, ssq_out=None, a_scale: float = 0.0, out_scale: float = 0.0) -> None:
    g = _lib.GemmArgs()
    g.M, g.N, g.K, g.precision = M, N, K, precision
    g.A, g.W, g.bias = A.data_ptr(), W.data_ptr(), _lib.ptr(bias)
    g.ssq_in = _lib.ptr(ssq_in)
    g.ssq_parts = 0 if ssq_in is None else int(ssq_in.shape[-1])
    g.rms_dim, g.rms_eps = rms_dim, rms_eps
    g.resid, g.act = _lib.ptr(resid), act
    g.out_f32, g.out_bf16, g.out_precision = _lib.ptr(out_f32), _lib.ptr(out_bf16), out_precision
    g.ssq_out = _lib.ptr(ssq_out)
    g.a_scale, g.out_scale = a_scale, out_scale
    _lib.check(_lib.load().avd_gemm(C.byref(g), _lib.stream_ptr()))


def core_workspace(M: int, d: int, d_mlp: int, precision: int, device):
    """Caller-allocated buffers of avd_mmdt_blocks (include/avdiff_b200.h: avd_core_workspace)."""
    S = _lib.nseg(precision)
    bufs = {
        "x": torch.empty(M, d, device=device, dtype=torch.float32),
        "xa": torch.empty(S, M, d, device=device, dtype=torch.bfloat16),
        "ssq": torch.empty(M, d // 64, device=device, dtype=torch.float32),
        "qkv": torch.empty(M, 3 * d, device=device,
                           dtype=torch.bfloat16 if precision in (_lib.AVD_PREC_BF16, _lib.AVD_PREC_F16)
                           else torch.float32),   # 16-bit storage (bf16 or fp16 bits, per the precision mode)
        "att": torch.empty(S, M, d, device=device, dtype=torch.bfloat16),
        "hmid": torch.empty(S, M, d_mlp, device=device, dtype=torch.bfloat16),
        # row-group counters of the chained out-proj -> fc1 -> fc2 -> QKV launch (zero between launches)
        "chain_flags": torch.zeros(_lib.AVD_CHAIN_FLAG_BYTES // 4, device=device, dtype=torch.int32),
    }
    ws = _lib.CoreWorkspace()
    for k, t in bufs.items():
        setattr(ws, k, t.data_ptr())
    ws.key_mask = None
    return ws, bufs


def pack_key_padding_mask(mask: torch.Tensor) -> torch.Tensor:
    """[B, N] bool (True = padding key, nn.MultiheadAttention's convention) -> int32 [B, ceil(N/32)] bit words, the
    layout avd_core_workspace.key_mask documents."""
    B, N = mask.shape
    words = (N + 31) // 32
    m = torch.zeros(B, words * 32, device=mask.device, dtype=torch.int64)
    m[:, :N] = mask.to(torch.int64)
    weights = torch.ones(32, device=mask.device, dtype=torch.int64) << torch.arange(32, device=mask.device)
    packed = (m.view(B, words, 32) * weights).sum(-1)                  # 0 .. 2^32-1
    return torch.where(packed >= 2 ** 31, packed - 2 ** 32, packed).to(torch.int32).contiguous()
