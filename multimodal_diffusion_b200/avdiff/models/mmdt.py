"""mmdt.py -- the MMDiT denoiser core on sm_100a (mirror of avdiff/models/mmdt.py).

Same classes, constructor arguments, parameter names (``state_dict`` keys) and construction order as the
reference, so ``torch.manual_seed(s); MMDiT(**cfg)`` draws the same random init and
``load_state_dict(reference_module.state_dict())`` works.  The parameters are fp32 masters; ``forward`` packs
them once (bf16 hi/mid/lo segments, RMSNorm scales folded into the consuming Linear) and runs the blocks
through the C ABI (avd_mmdt_blocks: tcgen05 GEMMs with fused RMSNorm-prologue / bias / GELU / residual
epilogues + the joint-attention kernel).  Eval / no_grad only: this is the sampling path.

``precision``: "fp32" (6-pass split-bf16 tensor-core GEMMs + fp32 attention; rel-L2 ~1e-6 vs the reference),
"bf16x3", "bf16", or "f16" (the mode bench.py times: fp16 operands at the same tensor-core rate as bf16 with 3 more
mantissa bits; the residual-stream operand carries a fixed 2^-8 scale so that it cannot overflow).  Default comes from $AVDIFF_B200_PRECISION, else "fp32".
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn

from ... import _lib, packing


def default_precision() -> str:
    return os.environ.get("AVDIFF_B200_PRECISION", "fp32")


def _no_grad_only(mod: nn.Module) -> None:
    if torch.is_grad_enabled() and any(p.requires_grad for p in mod.parameters()):
        raise RuntimeError("multimodal_diffusion_b200 implements the sampling (inference) path only: call under "
                           "torch.no_grad() (the reference's sampler is @torch.no_grad, sample_clip.py:220)")
    if mod.training and any(isinstance(m, nn.Dropout) and m.p > 0.0 for m in mod.modules()):
        raise RuntimeError("module is in training mode with dropout > 0, which the reference would apply; the kernels "
                           "implement the eval-mode forward only: call .eval() first (sample_clip.py:98,108)")


class RMSNorm(nn.Module):
    """Reference: mmdt.py:33-42.  y = scale * x / (||x||_2 / sqrt(d) + eps) -- eps is added to the rms."""

    def __init__(self, d: int, eps: float = 1e-6):
        super().__init__()
        self.scale = nn.Parameter(torch.ones(d))
        self.eps = eps

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _no_grad_only(self)
        xc = _lib.require_cuda_f32(x, "x")
        d = xc.shape[-1]
        rows = xc.numel() // d
        out = torch.empty_like(xc)
        if rows > 0:
            _lib.check(_lib.load().avd_rmsnorm_rows(xc.data_ptr(), self.scale.data_ptr(), 1, rows, 0, rows, d,
                                                    float(self.eps), out.data_ptr(), None, 0, _lib.stream_ptr()))
        return out


def build_norm(kind: str, d: int) -> nn.Module:
    """Reference: mmdt.py:45-46."""
    return RMSNorm(d) if kind.lower() == "rmsnorm" else nn.LayerNorm(d)


class MHA(nn.Module):
    """Reference: mmdt.py:51-61.  Holds an nn.MultiheadAttention purely as the parameter container
    (in_proj_weight [3d,d] rows q;k;v, in_proj_bias, out_proj.{weight,bias}); forward runs
    QKV GEMM -> joint attention kernel -> out-proj GEMM."""

    def __init__(self, d_model: int, n_heads: int, attn_dropout: float = 0.0, resid_dropout: float = 0.0):
        super().__init__()
        self.mha = nn.MultiheadAttention(embed_dim=d_model, num_heads=n_heads, dropout=attn_dropout, batch_first=True)
        self.drop = nn.Dropout(resid_dropout)
        self.precision = default_precision()

    def forward(self, x: torch.Tensor, attn_mask: Optional[torch.Tensor] = None,
                key_padding_mask: Optional[torch.Tensor] = None):
        _no_grad_only(self)
        if attn_mask is not None:
            raise NotImplementedError("attn_mask is not used by MMDiT or the sampler (SURVEY.md §8 N4)")
        xc = _lib.require_cuda_f32(x, "x")
        B, N, d = xc.shape
        M, H = B * N, self.mha.num_heads
        bits = None
        if key_padding_mask is not None:
            if key_padding_mask.shape != (B, N) or key_padding_mask.dtype != torch.bool:
                raise ValueError("key_padding_mask must be a bool tensor of shape [B, N]")
            bits = packing.pack_key_padding_mask(key_padding_mask.to(xc.device))
        prec = _lib.precision_code(self.precision)
        wq = packing.pack_linear_weight(self.mha.in_proj_weight.detach(), prec)
        wo = packing.pack_linear_weight(self.mha.out_proj.weight.detach(), prec)
        xa, _ = packing.rows_prepare(xc.view(M, d), prec)
        S = _lib.nseg(prec)
        f16 = prec == _lib.AVD_PREC_F16
        tc = prec == _lib.AVD_PREC_BF16 or f16
        qkv = torch.empty(M, 3 * d, device=xc.device, dtype=torch.bfloat16 if tc else torch.float32)
        packing.gemm(xa, wq, M, 3 * d, d, prec, bias=self.mha.in_proj_bias.detach(),
                     out_bf16=qkv if tc else None, out_f32=None if tc else qkv,
                     out_precision=prec if f16 else _lib.AVD_PREC_BF16,
                     a_scale=1.0 / _lib.AVD_F16_XSCALE if f16 else 0.0)   # rows_prepare pre-scales fp16 operands
        att = torch.empty(S, M, d, device=xc.device, dtype=torch.bfloat16)
        _lib.check(_lib.load().avd_attention_masked(qkv.data_ptr(), 0 if tc else 1, att.data_ptr(), prec, B, N, H,
                                                    d // H, None if bits is None else bits.data_ptr(),
                                                    _lib.stream_ptr()))
        y = torch.empty(M, d, device=xc.device, dtype=torch.float32)
        packing.gemm(att, wo, M, d, d, prec, bias=self.mha.out_proj.bias.detach(), out_f32=y)
        return y.view(B, N, d)


class MLP(nn.Module):
    """Reference: mmdt.py:66-83.  fc1 -> GELU(erf) -> fc2; the GELU lives in fc1's GEMM epilogue."""

    def __init__(self, d_model: int, mlp_ratio: float = 4.0, dropout: float = 0.0):
        super().__init__()
        hidden = int(d_model * mlp_ratio)
        self.fc1 = nn.Linear(d_model, hidden)
        self.fc2 = nn.Linear(hidden, d_model)
        self.drop = nn.Dropout(dropout)
        nn.init.xavier_uniform_(self.fc1.weight); nn.init.zeros_(self.fc1.bias)
        nn.init.xavier_uniform_(self.fc2.weight); nn.init.zeros_(self.fc2.bias)
        self.precision = default_precision()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _no_grad_only(self)
        xc = _lib.require_cuda_f32(x, "x")
        d = xc.shape[-1]
        M = xc.numel() // d
        hid = self.fc1.out_features
        prec = _lib.precision_code(self.precision)
        w1 = packing.pack_linear_weight(self.fc1.weight.detach(), prec)
        w2 = packing.pack_linear_weight(self.fc2.weight.detach(), prec)
        xa, _ = packing.rows_prepare(xc.view(M, d), prec)
        S = _lib.nseg(prec)
        u = torch.empty(S, M, hid, device=xc.device, dtype=torch.bfloat16)
        # 16-bit modes take the epilogue's packed GELU exactly as the fused core does (avd_api.cu core_blocks: act = 2)
        act = 2 if prec in (_lib.AVD_PREC_BF16, _lib.AVD_PREC_F16) else 1
        packing.gemm(xa, w1, M, hid, d, prec, bias=self.fc1.bias.detach(), act=act, out_bf16=u, out_precision=prec,
                     a_scale=1.0 / _lib.AVD_F16_XSCALE if prec == _lib.AVD_PREC_F16 else 0.0)
        y = torch.empty(M, d, device=xc.device, dtype=torch.float32)
        packing.gemm(u, w2, M, d, hid, prec, bias=self.fc2.bias.detach(), out_f32=y)
        return y.view(xc.shape)


class _CoreRunner:
    """Packs a set of blocks once (re-packs when a parameter changes) and runs avd_mmdt_blocks."""

    def __init__(self):
        self._packed: Optional[packing.PackedCore] = None
        self._key = None

    def run(self, owner: nn.Module, sd, n_layers: int, n_heads: int, precision, x: torch.Tensor,
            apply_final_norm: bool, key_padding_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        xc = _lib.require_cuda_f32(x, "x")
        B, N, d = xc.shape
        prec = _lib.precision_code(precision)
        key = (prec, str(xc.device), tuple((p.data_ptr(), p._version) for p in owner.parameters()))
        if self._packed is None or self._key != key:
            self._packed = packing.PackedCore(sd, n_layers, n_heads, prec, xc.device)
            self._key = key
        pc = self._packed
        M = B * N
        out = torch.empty(B, N, d, device=xc.device, dtype=torch.float32)
        if M == 0:
            return out
        xa, ssq = packing.rows_prepare(xc.view(M, d), prec)
        ws, bufs = packing.core_workspace(M, d, pc.d_mlp, prec, xc.device)
        if key_padding_mask is not None:
            if key_padding_mask.shape != (B, N) or key_padding_mask.dtype != torch.bool:
                raise ValueError("key_padding_mask must be a bool tensor of shape [B, N]")
            bits = packing.pack_key_padding_mask(key_padding_mask.to(xc.device))
            ws.key_mask = bits.data_ptr()
        _lib.check(_lib.load().avd_mmdt_blocks(C.byref(pc.struct), xc.data_ptr(), xa.data_ptr(), ssq.data_ptr(),
                                               C.byref(ws), B, N, _lib.stream_ptr()))
        if not apply_final_norm:
            return bufs["x"].view(B, N, d)
        if pc.layernorm:
            _lib.check(_lib.load().avd_layernorm_act(bufs["x"].data_ptr(), pc.final_ln_w.data_ptr(),
                                                     pc.final_ln_b.data_ptr(), M, d, 1e-5, 0, out.data_ptr(), None, 0,
                                                     _lib.stream_ptr()))
            return out
        _lib.check(_lib.load().avd_rmsnorm_rows(bufs["x"].data_ptr(), pc.final_scale.data_ptr(), 1, M, 0, M, d,
                                                1e-6, out.data_ptr(), None, 0, _lib.stream_ptr()))
        return out


class Block(nn.Module):
    """Reference: mmdt.py:88-99.  x = x + MHA(norm1(x)); x = x + MLP(norm2(x)) (eval: dropouts are identity)."""

    def __init__(self, d_model: int, n_heads: int, mlp_ratio: float, dropout: float, attn_dropout: float, norm: str):
        super().__init__()
        self.norm1 = build_norm(norm, d_model)
        self.attn = MHA(d_model, n_heads, attn_dropout=attn_dropout, resid_dropout=dropout)
        self.norm2 = build_norm(norm, d_model)
        self.mlp = MLP(d_model, mlp_ratio=mlp_ratio, dropout=dropout)
        self.n_heads = n_heads
        self.precision = default_precision()
        self._runner = _CoreRunner()

    def forward(self, x: torch.Tensor, attn_mask: Optional[torch.Tensor] = None,
                key_padding_mask: Optional[torch.Tensor] = None):
        _no_grad_only(self)
        if attn_mask is not None:
            raise NotImplementedError("attn_mask is not used by MMDiT or the sampler (SURVEY.md §8 N4)")
        sd = {"blocks.0." + k: v for k, v in self.state_dict().items()}
        if isinstance(self.norm1, RMSNorm):
            sd["final_norm.scale"] = torch.ones_like(self.norm1.scale)
        else:                                                   # the runner only uses it to pick the norm kind / width
            sd["final_norm.weight"] = torch.ones_like(self.norm1.weight)
            sd["final_norm.bias"] = torch.zeros_like(self.norm1.bias)
        return self._runner.run(self, sd, 1, self.n_heads, self.precision, x, apply_final_norm=False,
                                key_padding_mask=key_padding_mask)


@dataclass
class MMDiTCfg:
    d_model: int = 1024
    n_layers: int = 16
    n_heads: int = 16
    mlp_ratio: float = 4.0
    dropout: float = 0.1
    attn_dropout: float = 0.0
    norm: str = "rmsnorm"
    rope: bool = False
    token_dropout: float = 0.0


class MMDiT(nn.Module):
    """Reference: mmdt.py:116-149.  Stack of pre-norm self-attention blocks over the joint
    [video ; audio] token sequence + final norm.  forward(x [B,N,d] fp32 CUDA, key_padding_mask=None)."""

    def __init__(self, d_model=1024, n_layers=16, n_heads=16, mlp_ratio=4.0, dropout=0.1, attn_dropout=0.0,
                 norm="rmsnorm", rope=False, token_dropout=0.0, precision: Optional[str] = None):
        super().__init__()
        self.cfg = MMDiTCfg(d_model, n_layers, n_heads, mlp_ratio, dropout, attn_dropout, norm, rope, token_dropout)
        self.blocks = nn.ModuleList([
            Block(d_model, n_heads, mlp_ratio, dropout, attn_dropout, norm) for _ in range(n_layers)
        ])
        self.final_norm = build_norm(norm, d_model)
        self.precision = precision or default_precision()
        self._runner = _CoreRunner()

    def forward(self, x: torch.Tensor, key_padding_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        _no_grad_only(self)
        if self.training and self.cfg.token_dropout > 0.0:
            raise NotImplementedError("token_dropout is a training-time feature; the sampler runs in eval mode")
        if self.cfg.d_model % self.cfg.n_heads or (self.cfg.d_model // self.cfg.n_heads) != 64 and \
                _lib.precision_code(self.precision) in (_lib.AVD_PREC_BF16, _lib.AVD_PREC_F16):
            raise NotImplementedError("the tcgen05 attention kernel is built for head_dim 64")
        return self._runner.run(self, self.state_dict(), self.cfg.n_layers, self.cfg.n_heads, self.precision, x,
                                apply_final_norm=True, key_padding_mask=key_padding_mask)
