"""Shared helpers of the test-suite: golden loading, seeded weights, the oracle import."""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
GOLDEN = ROOT / "tests" / "golden"
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))



class _LazyOracle:
    """`orc.<name>` resolves into oracle/avdiff_oracle.py on first use.  The oracle is imported lazily so that callers
    which only need `randn` / `seeded_modules` (bench.py's own arm, tools/) never load anything under oracle/: only the
    tests, `__graft_entry__.smoke()` and bench.py's CPU legs touch `orc`."""

    def __getattr__(self, name):
        from oracle import avdiff_oracle
        return getattr(avdiff_oracle, name)


orc = _LazyOracle()

SEEDS = {"adapt_v": 100, "adapt_a": 101, "core": 102, "head": 103}   # same as tests/golden/make_golden.py

MVP = dict(d_model=512, n_layers=8, n_heads=8, mlp_ratio=4.0, dropout=0.1, attn_dropout=0.0, norm="rmsnorm",
           rope=False, token_dropout=0.0)


def golden(name: str):
    return np.load(GOLDEN / f"{name}.npz", allow_pickle=False)


def randn(seed: int, *shape) -> torch.Tensor:
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed), dtype=torch.float32)


def seeded_modules():
    """The mirror modules built with the constructor sequence / seeds of make_golden.seeded_modules; their
    random init must equal the reference's (checked against the checksums stored in mvp_forward.npz)."""
    from multimodal_diffusion_b200.avdiff.models.infer.sample_clip import LinearAdapter
    from multimodal_diffusion_b200.avdiff.models.mmdt import MMDiT
    from multimodal_diffusion_b200.avdiff.models.heads.noise_heads import MultiModalNoiseHead
    torch.manual_seed(SEEDS["adapt_v"]); adapt_v = LinearAdapter(256, 256)
    torch.manual_seed(SEEDS["adapt_a"]); adapt_a = LinearAdapter(32, 256)
    torch.manual_seed(SEEDS["core"]); core = MMDiT(**MVP).eval()
    torch.manual_seed(SEEDS["head"])
    head = MultiModalNoiseHead(input_dims={"video": 512, "audio": 512}, output_dims={"video": 256, "audio": 32},
                               hidden_dim=512, num_shared_layers=2, num_modality_specific_layers=1, dropout=0.1,
                               activation="gelu").eval()
    return adapt_v, adapt_a, core, head


def weights_dict(mods=None):
    adapt_v, adapt_a, core, head = mods or seeded_modules()
    return {"adapt_v": adapt_v.state_dict(), "adapt_a": adapt_a.state_dict(), "core": core.state_dict(),
            "head": head.state_dict()}


def rel_l2(a, b) -> float:
    a = torch.as_tensor(np.asarray(a.detach().cpu() if isinstance(a, torch.Tensor) else a)).double()
    b = torch.as_tensor(np.asarray(b.detach().cpu() if isinstance(b, torch.Tensor) else b)).double()
    den = float(b.norm())
    return float((a - b).norm()) / (den if den > 0 else 1.0)
