"""noise_heads.py -- epsilon-prediction heads on sm_100a (mirror of avdiff/models/heads/noise_heads.py).

Same classes, constructor arguments and ``state_dict`` keys as the reference (``input_proj.{m}``,
``shared.{i}.{0,1}``, ``spec.{m}``, ``out_proj.{m}``).  ``forward`` runs each present modality through
avd_head_forward: tcgen05 GEMMs (bias fused) + one LayerNorm+GELU(erf) row kernel per shared block.
Only the configuration the sampler builds is implemented on the GPU (sample_clip.py:99-107: shared trunk of
N blocks, identity modality-specific trunk, GELU); anything else raises NotImplementedError.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Union

import torch
import torch.nn as nn

from .... import _lib, packing
from ..mmdt import _no_grad_only, default_precision

__all__ = ["NoisePredictionHead", "MultiModalNoiseHead"]


def _make_activation(name: str) -> nn.Module:
    name = (name or "gelu").lower()
    if name == "gelu":
        return nn.GELU()
    if name == "relu":
        return nn.ReLU()
    if name == "leaky_relu":
        return nn.LeakyReLU(0.1)
    raise ValueError(f"Unsupported activation: {name}")


def _init_linear(m: nn.Module):
    if isinstance(m, nn.Linear):
        nn.init.xavier_uniform_(m.weight)
        if m.bias is not None:
            nn.init.zeros_(m.bias)


def _run_head(ph: packing.PackedHead, x2d: torch.Tensor) -> torch.Tensor:
    """x2d [M, d_in] fp32 -> eps [M, d_out] fp32 through avd_head_forward."""
    M = x2d.shape[0]
    S = _lib.nseg(ph.precision)
    dev = x2d.device
    eps = torch.empty(M, ph.d_out, device=dev, dtype=torch.float32)
    if M == 0:
        return eps
    h, _ = packing.rows_prepare(x2d, ph.precision)
    act_a = torch.empty(S, M, ph.hidden, device=dev, dtype=torch.bfloat16)
    act_b = torch.empty(S, M, ph.hidden, device=dev, dtype=torch.bfloat16)
    y = torch.empty(M, ph.hidden, device=dev, dtype=torch.float32)
    _lib.check(_lib.load().avd_head_forward(C.byref(ph.struct), h.data_ptr(), M, act_a.data_ptr(), act_b.data_ptr(),
                                            y.data_ptr(), eps.data_ptr(), _lib.stream_ptr()))
    return eps


class _PackCache:
    def __init__(self):
        self.key = None
        self.packed: Dict[str, packing.PackedHead] = {}

    def get(self, owner: nn.Module, modality: str, precision, device, builder) -> packing.PackedHead:
        key = (_lib.precision_code(precision), str(device),
               tuple((p.data_ptr(), p._version) for p in owner.parameters()))
        if key != self.key:
            self.key, self.packed = key, {}
        if modality not in self.packed:
            self.packed[modality] = builder()
        return self.packed[modality]


class NoisePredictionHead(nn.Module):
    """Reference: noise_heads.py:45-89.  Single-modality MLP head: Linear -> [LN, act, drop, Linear]*."""

    def __init__(self, input_dim: int, output_dim: int, hidden_dim: Optional[int] = None, num_layers: int = 2,
                 dropout: float = 0.0, activation: str = "gelu"):
        super().__init__()
        self.input_dim = int(input_dim)
        self.output_dim = int(output_dim)
        hidden_dim = int(hidden_dim or input_dim)
        self._activation = (activation or "gelu").lower()
        act = _make_activation(activation)
        layers = []
        if num_layers <= 1:
            layers.append(nn.Linear(self.input_dim, self.output_dim))
        else:
            layers += [nn.Linear(self.input_dim, hidden_dim), nn.LayerNorm(hidden_dim), act, nn.Dropout(dropout)]
            for _ in range(max(0, num_layers - 2)):
                layers += [nn.Linear(hidden_dim, hidden_dim), nn.LayerNorm(hidden_dim), act, nn.Dropout(dropout)]
            layers += [nn.Linear(hidden_dim, self.output_dim)]
        self.mlp = nn.Sequential(*layers)
        self.apply(_init_linear)
        self.precision = default_precision()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _no_grad_only(self)
        act_code = packing.activation_code(self._activation)
        xc = _lib.require_cuda_f32(x, "x")
        in_shape = xc.shape
        cur = xc.reshape(-1, in_shape[-1])
        M = cur.shape[0]
        prec = _lib.precision_code(self.precision)
        S = _lib.nseg(prec)
        mods = list(self.mlp)
        i = 0
        a, _ = packing.rows_prepare(cur, prec)
        out = None
        while i < len(mods):
            lin = mods[i]
            W = packing.pack_linear_weight(lin.weight.detach(), prec)
            N, K = lin.weight.shape
            y = torch.empty(M, N, device=xc.device, dtype=torch.float32)
            packing.gemm(a, W, M, N, K, prec, bias=lin.bias.detach(), out_f32=y)
            if i + 1 < len(mods) and isinstance(mods[i + 1], nn.LayerNorm):
                ln = mods[i + 1]
                a = torch.empty(S, M, N, device=xc.device, dtype=torch.bfloat16)
                _lib.check(_lib.load().avd_layernorm_act(y.data_ptr(), ln.weight.data_ptr(), ln.bias.data_ptr(), M, N,
                                                         float(ln.eps), act_code, None, a.data_ptr(), prec,
                                                         _lib.stream_ptr()))
                i += 4
            else:
                out = y
                i += 1
        return out.reshape(*in_shape[:-1], self.output_dim)


class MultiModalNoiseHead(nn.Module):
    """Reference: noise_heads.py:94-229.  forward(inputs: {modality: [..., d_in]}, return_dict=True)."""

    def __init__(self, input_dims: Dict[str, int], output_dims: Dict[str, int], hidden_dim: int = 512,
                 num_shared_layers: int = 2, num_modality_specific_layers: int = 1, dropout: float = 0.1,
                 activation: str = "gelu", share_parameters: bool = False, precision: Optional[str] = None):
        super().__init__()
        self.modalities = list(input_dims.keys())
        self.input_dims = {k: int(v) for k, v in input_dims.items()}
        self.output_dims = {k: int(v) for k, v in output_dims.items()}
        self.hidden_dim = int(hidden_dim)
        self.share_parameters = bool(share_parameters)
        self._activation = (activation or "gelu").lower()
        self._n_spec = int(num_modality_specific_layers)
        act = _make_activation(activation)

        self.input_proj = nn.ModuleDict({m: nn.Linear(self.input_dims[m], self.hidden_dim) for m in self.modalities})

        def _trunk_block():
            return nn.Sequential(nn.Linear(self.hidden_dim, self.hidden_dim), nn.LayerNorm(self.hidden_dim), act,
                                 nn.Dropout(dropout))

        self.shared = None if num_shared_layers <= 0 else \
            nn.Sequential(*[_trunk_block() for _ in range(num_shared_layers)])

        if num_modality_specific_layers <= 0:
            self.spec = nn.ModuleDict({m: nn.Identity() for m in self.modalities})
        elif self.share_parameters:
            self.shared_specific_trunk = \
                nn.Sequential(*[_trunk_block() for _ in range(num_modality_specific_layers - 1)]) \
                if num_modality_specific_layers > 1 else nn.Identity()
            self.spec = nn.ModuleDict({m: nn.Identity() for m in self.modalities})
        else:
            self.spec = nn.ModuleDict({
                m: (nn.Sequential(*[_trunk_block() for _ in range(num_modality_specific_layers - 1)])
                    if num_modality_specific_layers > 1 else nn.Identity())
                for m in self.modalities})

        self.out_proj = nn.ModuleDict({m: nn.Linear(self.hidden_dim, self.output_dims[m]) for m in self.modalities})
        self.apply(_init_linear)
        self.precision = precision or default_precision()
        self._cache = _PackCache()

    def get_output_dim(self, modality: str) -> int:
        return int(self.output_dims[modality])

    def ln_eps(self) -> float:
        """eps of the trunk's nn.LayerNorm modules (all built alike, noise_heads.py:139-146)."""
        for mod in self.modules():
            if isinstance(mod, nn.LayerNorm):
                return float(mod.eps)
        return 1e-5

    def forward(self, inputs: Dict[str, torch.Tensor], return_dict: bool = True) \
            -> Union[Dict[str, torch.Tensor], torch.Tensor]:
        _no_grad_only(self)
        outputs: Dict[str, torch.Tensor] = {}
        for m in self.modalities:
            if m not in inputs or inputs[m] is None:
                continue
            x = _lib.require_cuda_f32(inputs[m], f"inputs[{m!r}]")
            in_shape = x.shape
            sd = self.state_dict()
            ph = self._cache.get(self, m, self.precision, x.device,
                                 lambda: packing.PackedHead(sd, m, self.precision, x.device,
                                                            ln_eps=self.ln_eps(), activation=self._activation))
            eps = _run_head(ph, x.reshape(-1, in_shape[-1]))
            outputs[m] = eps.view(*in_shape[:-1], self.output_dims[m])
        if return_dict:
            return outputs
        for m in self.modalities:
            if m in outputs:
                return outputs[m]
        raise ValueError("No modalities found in inputs for MultiModalNoiseHead.forward")
