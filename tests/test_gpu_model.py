"""GPU: the MMDT core, the noise head and the fused DDIM+CFG loop through the mirror modules / FusedSampler,
against fixtures produced by the unmodified reference (tests/golden) and against the CPU oracle.

Tolerances are BASELINE.json's: per-step eps rel-L2 <= 1e-5 (fp32 mode) / <= 1e-2 (bf16 mode);
final 50-step latents rel-L2 <= 2e-2."""
import numpy as np
import pytest
import torch

from _common import golden, orc, randn, rel_l2, seeded_modules, weights_dict

pytestmark = pytest.mark.gpu

TOL_EPS_FP32 = 1e-5
TOL_EPS_BF16 = 1e-2
TOL_FINAL = 2e-2


@pytest.fixture(scope="module")
def dev(lib):
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def mods(dev):
    return tuple(m.to(dev) for m in seeded_modules())


@pytest.fixture(scope="module")
def w_cpu():
    return weights_dict()


def T(a):
    return torch.from_numpy(np.asarray(a))


@pytest.mark.parametrize("precision,tol", [("fp32", TOL_EPS_FP32), ("bf16x3", 2e-4), ("bf16", 8e-3), ("f16", 1.5e-3)])
def test_mmdit_forward_vs_reference(dev, mods, precision, tol):
    g = golden("mvp_forward")
    core = mods[2]
    core.precision = precision
    X = T(g["X"]).to(dev)
    with torch.no_grad():
        h = core(X)
        X_null = torch.cat([torch.zeros_like(X[:, :96]), X[:, 96:]], dim=1)
        hn = core(X_null)
    err_c, err_n = rel_l2(h, g["h_cond"]), rel_l2(hn, g["h_null"])
    print(f"[mmdit {precision}] rel-L2 cond {err_c:.3e} null {err_n:.3e}")
    assert err_c < tol and err_n < tol


def test_mmdit_forward_cta_pair_kernels(dev, mods):
    """The specialised epilogues of the cta_group::2 GEMM variant (option gemm_pair = 1) through a whole forward."""
    from multimodal_diffusion_b200 import _lib
    g = golden("mvp_forward")
    core = mods[2]
    core.precision = "f16"
    X = T(g["X"]).to(dev).repeat(8, 1, 1)        # 16 sequences: the row-tile count the sampling batch has
    with torch.no_grad():
        with _lib.option("gemm_pair", 0):
            base = core(X)
        with _lib.option("gemm_pair", 1):
            h = core(X)
    assert rel_l2(h[:2], g["h_cond"]) < 1.5e-3
    assert torch.equal(h, base)


def test_block_and_submodules_vs_reference(dev, mods, w_cpu):
    g = golden("mvp_forward")
    core = mods[2]
    X = T(g["X"]).to(dev)
    blk = core.blocks[0]
    blk.precision = "fp32"
    with torch.no_grad():
        out = blk(X)
        assert rel_l2(out, g["block0_out"]) < TOL_EPS_FP32
        blk.mlp.precision = blk.attn.precision = "fp32"
        sd = w_cpu["core"]
        hcpu = orc.rmsnorm(T(g["X"]), sd["blocks.0.norm1.scale"])
        want_attn = orc.mha(hcpu, sd["blocks.0.attn.mha.in_proj_weight"], sd["blocks.0.attn.mha.in_proj_bias"],
                            sd["blocks.0.attn.mha.out_proj.weight"], sd["blocks.0.attn.mha.out_proj.bias"], 8)
        assert rel_l2(blk.attn(blk.norm1(X)), want_attn) < TOL_EPS_FP32
        want_mlp = orc.gelu_erf(hcpu @ sd["blocks.0.mlp.fc1.weight"].t() + sd["blocks.0.mlp.fc1.bias"]) \
            @ sd["blocks.0.mlp.fc2.weight"].t() + sd["blocks.0.mlp.fc2.bias"]
        assert rel_l2(blk.mlp(blk.norm1(X)), want_mlp) < TOL_EPS_FP32


@pytest.mark.parametrize("precision,tol", [("fp32", TOL_EPS_FP32), ("bf16x3", 1e-4), ("f16", 2e-3),
                                           ("bf16", TOL_EPS_BF16)])
def test_head_vs_reference(dev, mods, precision, tol):
    g = golden("mvp_forward")
    head = mods[3]
    head.precision = precision
    h_c, h_n = T(g["h_cond"]).to(dev), T(g["h_null"]).to(dev)
    with torch.no_grad():
        e_c = head({"video": h_c[:, :96], "audio": h_c[:, 96:]})
        e_n = head({"video": None, "audio": h_n[:, 96:]})
        first = head({"audio": h_c[:, 96:]}, return_dict=False)
    assert set(e_c) == {"video", "audio"} and set(e_n) == {"audio"}
    assert e_c["audio"].shape == (2, 37, 32) and e_c["video"].shape == (2, 96, 256)
    errs = (rel_l2(e_c["audio"], g["eps_cond_audio"]), rel_l2(e_n["audio"], g["eps_null_audio"]),
            rel_l2(e_c["video"], g["eps_cond_video"]))
    print(f"[head {precision}] rel-L2 {errs}")
    assert max(errs) < tol and torch.equal(first, e_c["audio"])
    with pytest.raises(ValueError), torch.no_grad():
        head({}, return_dict=False)                      # noise_heads.py:229


def test_noise_prediction_head_and_adapters(dev, mods, w_cpu):
    from multimodal_diffusion_b200.avdiff.models.heads.noise_heads import NoisePredictionHead
    from multimodal_diffusion_b200.avdiff.models.infer.sample_clip import add_sinusoidal_timestep
    g = golden("mvp_forward")
    torch.manual_seed(7)
    nph = NoisePredictionHead(512, 256, hidden_dim=512, num_layers=3).eval().to(dev)
    with torch.no_grad():
        out = nph(T(g["h_cond"]).to(dev)[:, :13])
    assert out.shape == (2, 13, 256) and rel_l2(out, g["nph_out"]) < TOL_EPS_FP32
    adapt_v, adapt_a = mods[0], mods[1]
    tok_v, tok_a = randn(1, 2, 96, 256), randn(2, 2, 37, 32)
    with torch.no_grad():
        xv, xa = adapt_v(tok_v.to(dev)), adapt_a(tok_a.to(dev))
    assert rel_l2(xv, orc.linear_adapter(w_cpu["adapt_v"], tok_v)) < 2e-6
    assert rel_l2(xa, orc.linear_adapter(w_cpu["adapt_a"], tok_a)) < 2e-6
    cat = add_sinusoidal_timestep(xa, torch.full((2,), 999), 256)
    assert cat.shape == (2, 37, 512)
    assert float((cat[..., 256:].cpu() - orc.timestep_embedding(torch.full((2,), 999), 256)[:, None]).abs().max()) < 2e-4


def _sampler(mods, target, B, core_precision, head_precision, steps=50):
    from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig
    cfg = HotPathConfig(sampler_steps=steps, guidance=3.5)
    return FusedSampler(cfg, mods[0], mods[1], mods[2], mods[3], target, B, device="cuda:0",
                        core_precision=core_precision, head_precision=head_precision)


@pytest.mark.parametrize("target", ["audio", "video"])
@pytest.mark.parametrize("core_p,head_p,tol", [("fp32", "fp32", TOL_EPS_FP32), ("bf16", "bf16x3", TOL_EPS_BF16),
                                               ("f16", "bf16x3", TOL_EPS_BF16), ("f16", "f16", TOL_EPS_BF16)])
def test_sampler_first_steps_vs_oracle(dev, mods, w_cpu, target, core_p, head_p, tol):
    """Per-step eps of both CFG branches, the CFG-combined latent-layout eps and z_prev, B = 2, steps 0 and 1."""
    B = 2
    fs = _sampler(mods, target, B, core_p, head_p)
    zp = randn(41, B, 8, 12, 16, 16) if target == "audio" else randn(41, B, 8, 150)
    zi = randn(42, B, 8, 150) if target == "audio" else randn(42, B, 8, 12, 16, 16)
    cfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5)
    rec = []
    with torch.no_grad():
        # drive the oracle for two steps only
        ab, ts = cfg.alpha_bar(), cfg.timesteps()
        z_or = zi.clone()
        fs.prepare(zp.to(dev))
        z_gpu = zi.to(dev).clone()
        tap = torch.empty(2 * B, fs.Nt, fs.Dt, device=dev)
        for i in range(2):
            t_now, t_prev = ts[i].repeat(B), ts[i + 1].repeat(B)
            z_v, z_a = (zp, z_or) if target == "audio" else (z_or, zp)
            e_c, e_n, e_hat = orc.eps_for_step(cfg, w_cpu, target, z_v, z_a, t_now)
            # feed the ORACLE's z_t to the GPU step so that each step is compared on identical inputs
            z_gpu.copy_(z_or)
            fs.step(i, z_gpu, tap)
            z_or = orc.ddim_step(z_or, t_now, t_prev, orc.untokenise(cfg, target, e_hat, z_or), ab)
            ec_g, en_g = tap[:B].cpu(), tap[B:].cpu()
            errs = (rel_l2(ec_g, e_c), rel_l2(en_g, e_n))
            e_hat_g = en_g + 3.5 * (ec_g - en_g)
            err_hat, err_z = rel_l2(e_hat_g, e_hat), rel_l2(z_gpu, z_or)
            print(f"[{target} {core_p}/{head_p} step {i}] eps cond/null {errs[0]:.2e}/{errs[1]:.2e} "
                  f"cfg {err_hat:.2e} z_prev {err_z:.2e}")
            assert max(errs) < tol, (i, errs)
            assert err_z < (2e-2 if core_p == "bf16" else 10 * tol)
            rec.append((err_hat, err_z))
    if core_p == "fp32":
        assert max(r[0] for r in rec) < 5 * tol
    if core_p == "f16":
        # the half-precision tolerance (1e-2) also holds for the CFG-COMBINED eps, which guidance 3.5 amplifies
        assert max(r[0] for r in rec) < tol


@pytest.mark.parametrize("name,target", [("v2a", "audio"), ("a2v", "video")])
@pytest.mark.parametrize("core_p,head_p,tol", [("fp32", "fp32", 2e-3), ("bf16", "bf16x3", TOL_FINAL),
                                               ("f16", "bf16x3", TOL_FINAL), ("f16", "f16", TOL_FINAL)])
def test_sampler_50_steps_vs_reference(dev, mods, name, target, core_p, head_p, tol):
    """Final latent of the reference's own sample_one_direction run (B = 1, 50 steps, injected latents)."""
    g = golden(f"sample_{name}")
    fs = _sampler(mods, target, 1, core_p, head_p)
    zp, zi = T(g["z_prompt"]).to(dev), T(g["z_init"]).to(dev)
    z_eager = fs.sample_eager(zp, zi)
    z_graph = fs.sample(zp, zi)
    err = rel_l2(z_eager, g["z_final"])
    print(f"[{name} {core_p}/{head_p}] final-latent rel-L2 {err:.3e}; |z| {float(z_eager.norm()):.4e} "
          f"(reference {float(g['z_norms'][-1]):.4e})")
    assert torch.equal(z_eager, z_graph)                  # the CUDA graph replays exactly the eager launches
    assert err < tol
    if target == "audio":
        # frames 148,149 see eps = 0 every step: a pure fp32 elementwise chain -> bit-exact
        assert np.array_equal(z_eager[..., 148:].cpu().numpy(), g["z_final"][..., 148:])


def test_full_size_v2a_batch8_properties(dev, mods):
    """BASELINE.json configs[1] (V2A, B = 8, 50 steps) through the graph path: size-independent properties."""
    B = 8
    fs = _sampler(mods, "audio", B, "bf16", "bf16x3")
    zp, zi = randn(1, B, 8, 12, 16, 16), randn(2, B, 8, 150)
    z = fs.sample(zp.to(dev), zi.to(dev))
    assert torch.isfinite(z).all()
    # (1) clip independence: clip b of the batch equals the same clip sampled alone (no cross-clip term anywhere)
    fs1 = _sampler(mods, "audio", 1, "bf16", "bf16x3")
    for b in (0, 5):
        z1 = fs1.sample(zp[b:b + 1].to(dev), zi[b:b + 1].to(dev))
        assert rel_l2(z[b:b + 1], z1) < 1e-5
    # (2) the un-modelled tail (frames 148,149) follows the eps = 0 DDIM recursion exactly (bit-exact chain)
    cfg = orc.HotPathConfig()
    ab, ts = cfg.alpha_bar(), cfg.timesteps()
    tail = zi[..., 148:].clone()
    for i in range(50):
        tail = orc.ddim_step(tail, ts[i].repeat(B), ts[i + 1].repeat(B), torch.zeros_like(tail), ab)
    assert torch.equal(z[..., 148:].cpu(), tail)
    # (3) determinism: replaying the graph on the same inputs is bitwise reproducible
    assert torch.equal(fs.sample(zp.to(dev), zi.to(dev)), z)


@pytest.mark.parametrize("core_p,head_p,tol", [("fp32", "fp32", 1e-5), ("f16", "f16", 5e-3)])
def test_full_size_a2v_batch64_sharding_invariance(dev, mods, core_p, head_p, tol):
    """BASELINE.json configs[2] (A2V, B = 64): the clip-sharded result (8 shards of 8, as 8 GPUs would compute it)
    equals the single-batch result clip by clip.  5 steps keep the test short; the property is step-independent.
    Batch 64 runs the persistent GEMMs (256-wide column-split tiles for QKV / fc1), batch 8 the one-tile kernels: the two
    accumulate the k-blocks in a different order, so outside fp32 mode the results agree to the rounding noise of the
    16-bit activations (amplified by guidance 3.5), not bit for bit."""
    from multimodal_diffusion_b200.sampler import clip_noise, shard_range
    n = 64
    zp = torch.stack([clip_noise((8, 150), i, 1000) for i in range(n)])
    zi = torch.stack([clip_noise((8, 12, 16, 16), i, 2000) for i in range(n)])
    full = _sampler(mods, "video", n, core_p, head_p, steps=5).sample(zp.to(dev), zi.to(dev))
    fs8 = _sampler(mods, "video", 8, core_p, head_p, steps=5)
    worst = 0.0
    for r in range(8):
        lo, hi = shard_range(n, r, 8)
        part = fs8.sample(zp[lo:hi].to(dev), zi[lo:hi].to(dev))
        worst = max(worst, rel_l2(part, full[lo:hi]))
    print(f"[sharding {core_p}] worst shard rel-L2 {worst:.3e}")
    assert worst < tol


def test_long_context_a2v_10s(dev, mods, w_cpu):
    """BASELINE.json configs[3]: 10 s clip, N = 320 + 125 = 445 tokens; two steps against the oracle."""
    from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig
    cfg = HotPathConfig(sampler_steps=50, guidance=3.5, Fa=500, video_latent=(40, 16, 16))
    assert cfg.Nv + cfg.Na == 445
    fs = FusedSampler(cfg, mods[0], mods[1], mods[2], mods[3], "video", 1, core_precision="fp32", head_precision="fp32")
    zp, zi = randn(51, 1, 8, 500), randn(52, 1, 8, 40, 16, 16)
    ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5, Fa=500)
    ab, ts = ocfg.alpha_bar(), ocfg.timesteps()
    with torch.no_grad():
        e_c, e_n, e_hat = orc.eps_for_step(ocfg, w_cpu, "video", zi, zp, ts[0].repeat(1))
        z_or = orc.ddim_step(zi, ts[0].repeat(1), ts[1].repeat(1), orc.untokenise(ocfg, "video", e_hat, zi), ab)
    fs.prepare(zp.to(dev))
    z = zi.to(dev).clone()
    tap = torch.empty(2, 320, 256, device=dev)
    fs.step(0, z, tap)
    assert rel_l2(tap[:1], e_c) < TOL_EPS_FP32 and rel_l2(tap[1:], e_n) < TOL_EPS_FP32
    assert rel_l2(z, z_or) < 1e-4


@pytest.mark.parametrize("precision,tol", [("fp32", TOL_EPS_FP32), ("f16", 2e-3)])
def test_mmdit_key_padding_mask(dev, mods, w_cpu, precision, tol):
    """MMDiT.forward(x, key_padding_mask) (mmdt.py:134-147; §8f N4) against the oracle; outputs at padded positions are
    defined too (padded QUERIES still attend to the valid keys), so the whole tensor is compared."""
    core = mods[2]
    core.precision = precision
    x = randn(81, 3, 133, 512)
    mask = torch.zeros(3, 133, dtype=torch.bool)
    mask[0, 120:] = True
    mask[1, 96:] = True                                  # e.g. a clip without audio tokens
    mask[2, 5:9] = True
    with torch.no_grad():
        got = core(x.to(dev), key_padding_mask=mask.to(dev)).cpu()
        want = orc.mmdit_forward(w_cpu["core"], x, 8, 8, key_padding_mask=mask)
        plain = core(x.to(dev)).cpu()
    assert rel_l2(got, want) < tol
    assert rel_l2(plain[1], want[1]) > 5 * tol           # the mask matters (37 of 133 keys padded)


@pytest.mark.parametrize("precision,tol", [("fp32", TOL_EPS_FP32), ("f16", 2e-3), ("bf16", 1e-2)])
def test_mmdit_layernorm_variant(dev, precision, tol):
    """MMDiT(norm="layernorm") (mmdt.py:45-46; §8f N4): LayerNorm runs as its own pass in front of the QKV / fc1 GEMMs."""
    from multimodal_diffusion_b200.avdiff.models.mmdt import MMDiT
    torch.manual_seed(5)
    core = MMDiT(d_model=512, n_layers=3, n_heads=8, mlp_ratio=4.0, dropout=0.0, norm="layernorm").eval()
    with torch.no_grad():                       # non-trivial affine parameters
        for k, v in core.state_dict().items():
            if "norm" in k:
                v.copy_(1.0 + 0.2 * randn(hash(k) % 1000, *v.shape) if k.endswith("weight") else 0.1 * randn(hash(k) % 999, *v.shape))
    sd = {k: v.clone() for k, v in core.state_dict().items()}
    x = randn(91, 2, 133, 512) * 2 + 0.7        # a row mean the norm has to remove
    core = core.to(dev)
    core.precision = precision
    with torch.no_grad():
        got = core(x.to(dev)).cpu()
        blk = core.blocks[1]
        blk.precision = precision
        got_blk = blk(x.to(dev)).cpu()
    assert rel_l2(got, orc.mmdit_forward(sd, x, 3, 8)) < tol
    assert rel_l2(got_blk, orc.block_forward(sd, 1, x, 8)) < tol


def test_stream_sample_windows_as_batch(dev, mods):
    """§8f N2: the windows of a long prompt run as the batch dimension of the fused sampler; the result must not depend
    on the batch size (5 windows through a batch-2 sampler, last call padded, against a batch-1 sampler)."""
    from multimodal_diffusion_b200.avdiff.models.infer.stream_infer import stream_sample
    zp = randn(71, 5, 8, 12, 16, 16).to(dev)
    zi = randn(72, 5, 8, 150).to(dev)
    a = stream_sample(_sampler(mods, "audio", 2, "f16", "f16", steps=4), zp, z_init_windows=zi)
    b = stream_sample(_sampler(mods, "audio", 1, "f16", "f16", steps=4), zp, z_init_windows=zi)
    assert a.shape == (5, 8, 150) and torch.equal(a, b)
    c = stream_sample(_sampler(mods, "audio", 2, "f16", "f16", steps=4), zp, base_seed=9)
    d = stream_sample(_sampler(mods, "audio", 1, "f16", "f16", steps=4), zp, base_seed=9)
    assert torch.equal(c, d) and not torch.equal(a, c)
