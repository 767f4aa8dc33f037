"""GPU, round 2: parity on the configurations that are TIMED (VERDICT r1 "next round" item 1) and the generality rows
(SURVEY §8f N3 / N4): the oracle comparisons run on the same seeded inputs; tolerances are BASELINE.json's
(per-step eps rel-L2 <= 1e-5 fp32 / <= 1e-2 16-bit, final 50-step latents <= 2e-2)."""
import ctypes as C
import json
import math

import numpy as np
import pytest
import torch

from _common import GOLDEN, golden, orc, randn, rel_l2, seeded_modules, weights_dict

pytestmark = pytest.mark.gpu

TOL_EPS_FP32 = 1e-5
TOL_EPS_16 = 1e-2
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


def _sampler(mods, target, B, core_p="f16", head_p="f16", steps=50, **kw):
    from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig
    cfg_kw = {k: kw.pop(k) for k in ("Fa", "video_latent", "eta") if k in kw}
    cfg = HotPathConfig(sampler_steps=steps, guidance=3.5, **cfg_kw)
    return FusedSampler(cfg, mods[0], mods[1], mods[2], mods[3], target, B, device="cuda:0",
                        core_precision=core_p, head_precision=head_p, **kw)


def _step_vs_oracle(fs, ocfg, w, target, z_t, z_prompt, i, dev):
    """One GPU step fed the ORACLE's z_t; returns (err eps cond, err eps null, err CFG-combined eps, err z_prev)."""
    B = z_t.shape[0]
    ab, ts = ocfg.alpha_bar(), ocfg.timesteps()
    t_now, t_prev = ts[i].repeat(B), ts[i + 1].repeat(B)
    z_v, z_a = (z_prompt, z_t) if target == "audio" else (z_t, z_prompt)
    with torch.no_grad():
        e_c, e_n, e_hat = orc.eps_for_step(ocfg, w, target, z_v, z_a, t_now)
        z_or = orc.ddim_step(z_t, t_now, t_prev, orc.untokenise(ocfg, target, e_hat, z_t), ab)
    z_gpu = z_t.to(dev).clone()
    tap = torch.empty(2 * B, fs.Nt, fs.Dt, device=dev)
    fs.step(i, z_gpu, tap)
    ec_g, en_g = tap[:B].cpu(), tap[B:].cpu()
    e_hat_g = en_g + ocfg.guidance * (ec_g - en_g)
    return rel_l2(ec_g, e_c), rel_l2(en_g, e_n), rel_l2(e_hat_g, e_hat), rel_l2(z_gpu, z_or)


# ------------------------------------------------------------------------------------------ 1a: the timed configuration
def test_full_size_v2a_batch8_f16_graph_vs_oracle(dev, mods, w_cpu):
    """BASELINE.json configs[1] exactly as bench.py times it: V2A, B = 8, f16/f16, CUDA-graph path, 50 steps, against
    the oracle's 50-step loop; per-step eps for steps 0 / 1 / 49 per branch AND CFG-combined on the oracle's z_t."""
    B = 8
    fs = _sampler(mods, "audio", B)
    zp, zi = randn(1, B, 8, 12, 16, 16), randn(2, B, 8, 150)
    ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5)
    rec = []
    z_or = orc.sample_loop(ocfg, w_cpu, "audio", zp, zi, record=rec, record_steps=(0, 1, 49))
    z = fs.sample(zp.to(dev), zi.to(dev))
    err_final = rel_l2(z, z_or)
    print(f"[v2a B=8 f16 graph] final-latent rel-L2 {err_final:.3e}")
    assert err_final < TOL_FINAL
    assert torch.equal(fs.sample(zp.to(dev), zi.to(dev)), z)            # replay is deterministic
    fs.prepare(zp.to(dev))
    for r in rec:
        ec, en, ehat, ez = _step_vs_oracle(fs, ocfg, w_cpu, "audio", r["z_t"], zp, r["step"], dev)
        print(f"[v2a B=8 f16 step {r['step']}] eps cond/null {ec:.2e}/{en:.2e} cfg {ehat:.2e} z_prev {ez:.2e}")
        assert max(ec, en) < TOL_EPS_16 and ehat < TOL_EPS_16
    fs.check_errors()


# ------------------------------------------------------------------------------------------ 1b: persistent GEMM, directly
def _ps_case(dev, M, N, K, mode, bn=0):
    """avd_gemm forced onto gemm_ps_kernel (option gemm_v = 4) against an fp64 matmul of the rounded bf16 operands
    (computed with torch on the GPU: checker only).  mode: 'f32' generic, 'b16' 16-bit only (MODE 1), 'gelu' packed GELU
    (MODE 2), 'resid' residual + fp32 + 16-bit + ssq (MODE 3)."""
    from multimodal_diffusion_b200 import _lib, packing
    prec = _lib.AVD_PREC_BF16
    g = torch.Generator(device="cpu").manual_seed(M + N + K)
    A = torch.randn(M, K, generator=g).to(dev)
    W = (torch.randn(N, K, generator=g) / math.sqrt(K)).to(dev)
    bias = torch.randn(N, generator=g).to(dev)
    xa, _ = packing.rows_prepare(A, prec)
    Wp = packing.pack_linear_weight(W, prec)
    want = A.bfloat16().double() @ W.bfloat16().double().t() + bias.double()
    out_f32 = out_b = ssq = resid = None
    kw = {}
    if mode == "f32":
        out_f32 = torch.zeros(M, N, device=dev)
    elif mode == "b16":
        out_b = torch.zeros(1, M, N, device=dev, dtype=torch.bfloat16)
    elif mode == "gelu":
        out_b = torch.zeros(1, M, N, device=dev, dtype=torch.bfloat16)
        kw["act"] = 2
        want = 0.5 * want * (1.0 + torch.erf(want / math.sqrt(2.0)))
    else:
        resid = torch.randn(M, N, generator=g).to(dev)
        out_f32 = torch.zeros(M, N, device=dev)
        out_b = torch.zeros(1, M, N, device=dev, dtype=torch.bfloat16)
        ssq = torch.zeros(M, N // 64, device=dev)
        want = want + resid.double()
    with _lib.option("gemm_v", 4), _lib.option("gemm_bn", bn):
        packing.gemm(xa, Wp, M, N, K, prec, bias=bias, resid=resid, out_f32=out_f32, out_bf16=out_b,
                     out_precision=prec if out_b is not None else 0, ssq_out=ssq, **kw)
    torch.cuda.synchronize()
    if out_f32 is not None:
        assert rel_l2(out_f32, want) < 2e-6, (M, N, K, mode, bn)          # exact products, fp32 accumulation
    if out_b is not None:
        # one bf16 rounding of the result (2^-9 relative), plus the tanh-form GELU (<= 4.8e-4 absolute) in 'gelu'
        assert rel_l2(out_b[0].double(), want) < (4e-3 if mode == "gelu" else 2.5e-3), (M, N, K, mode, bn)
        worst = float((out_b[0].double() - want).abs().max() / want.abs().max())
        assert worst < 6e-3, (M, N, K, mode, bn, worst)
    if ssq is not None:
        assert rel_l2(ssq.sum(-1).double(), (want ** 2).sum(-1)) < 1e-5


@pytest.mark.parametrize("M", [4256, 17024, 32768])
@pytest.mark.parametrize("N,K,mode", [(1536, 512, "b16"), (512, 512, "resid"), (2048, 512, "gelu"), (512, 2048, "resid"),
                                      (1536, 512, "f32"), (512, 2048, "f32")])
def test_gemm_persistent_kernel_direct(dev, lib, M, N, K, mode):
    """gemm_ps_kernel through avd_gemm at the row counts of A2V batch 16 / 64 and the sweep, with the tile width its
    model picks (256-wide column-split tiles for the 16-bit-only GEMMs, 128-wide for the residual ones)."""
    _ps_case(dev, M, N, K, mode)
    out = (C.c_uint * 256)()
    assert lib.avd_debug_gemm_timeout(out) == 0 and out[0] == 0, "a barrier wait of the persistent GEMM timed out"
    assert lib.avd_check_errors() == 0


@pytest.mark.parametrize("bn", [256, 128, 64])
@pytest.mark.parametrize("N,K,mode", [(1536, 512, "b16"), (2048, 512, "gelu"), (512, 512, "resid"), (512, 2048, "f32")])
def test_gemm_persistent_every_tile_width(dev, lib, bn, N, K, mode):
    """Every instantiation <BN, MODE> the launcher can reach (forced tile width; combinations whose staging does not fit
    fall back to the model's choice), a ragged last row tile (M = 4256 + 37), and fewer tiles than SMs (M = 300)."""
    for M in (4293, 300):
        _ps_case(dev, M, N, K, mode, bn=bn)
    out = (C.c_uint * 256)()
    assert lib.avd_debug_gemm_timeout(out) == 0 and out[0] == 0


# ------------------------------------------------------------------------------------------ 1c: A2V batch 64, long context
def test_a2v_batch64_f16_first_steps_vs_oracle(dev, mods, w_cpu):
    """BASELINE.json configs[2]: A2V, B = 64 (persistent GEMMs, wide build_x) in the timed f16/f16 mode, steps 0 and 1."""
    B = 64
    fs = _sampler(mods, "video", B)
    zp, zi = randn(61, B, 8, 150), randn(62, B, 8, 12, 16, 16)
    ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5)
    ab, ts = ocfg.alpha_bar(), ocfg.timesteps()
    fs.prepare(zp.to(dev))
    z_or = zi.clone()
    for i in range(2):
        ec, en, ehat, ez = _step_vs_oracle(fs, ocfg, w_cpu, "video", z_or, zp, i, dev)
        print(f"[a2v B=64 f16 step {i}] eps cond/null {ec:.2e}/{en:.2e} cfg {ehat:.2e} z_prev {ez:.2e}")
        assert max(ec, en) < TOL_EPS_16 and ehat < TOL_EPS_16
        with torch.no_grad():
            t_now, t_prev = ts[i].repeat(B), ts[i + 1].repeat(B)
            _, _, e_hat = orc.eps_for_step(ocfg, w_cpu, "video", z_or, zp, t_now)
            z_or = orc.ddim_step(z_or, t_now, t_prev, orc.untokenise(ocfg, "video", e_hat, z_or), ab)
    fs.check_errors()


def test_long_context_f16_batch2_vs_oracle(dev, mods, w_cpu):
    """BASELINE.json configs[3]: 10 s clip, N = 445 joint tokens (multi-block attention, F16 template), B = 2, f16/f16."""
    B = 2
    fs = _sampler(mods, "video", B, Fa=500, video_latent=(40, 16, 16))
    zp, zi = randn(51, B, 8, 500), randn(52, B, 8, 40, 16, 16)
    ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5, Fa=500)
    fs.prepare(zp.to(dev))
    ec, en, ehat, ez = _step_vs_oracle(fs, ocfg, w_cpu, "video", zi, zp, 0, dev)
    print(f"[a2v 10 s B=2 f16 step 0] eps cond/null {ec:.2e}/{en:.2e} cfg {ehat:.2e} z_prev {ez:.2e}")
    assert max(ec, en) < TOL_EPS_16 and ehat < TOL_EPS_16


def _attn_check(dev, lib, n_seq, N, f16, tol, key_ramp=0.0):
    from multimodal_diffusion_b200 import _lib
    H, hd = 8, 64
    dt = torch.float16 if f16 else torch.bfloat16
    qkv = randn(3, n_seq * N, 3 * H * hd) * 1.5
    if key_ramp:          # scores rise with the key index by key_ramp / 2 nats over the sequence (every head's dim 0)
        v5 = qkv.view(n_seq, N, 3, H, hd)
        v5[:, :, 0, :, 0] = 4.0
        v5[:, :, 1, :, 0] = torch.linspace(-key_ramp, key_ramp, N).view(1, N, 1)
    qkv = qkv.to(dt).to(dev)
    # guard bands of 32 rows either side of the output (compute-sanitizer is closed on the pool): a store past the last
    # query row of a ragged tile, or before the first, would change them
    G = 32
    buf = torch.full((n_seq * N + 2 * G, H * hd), 7.0, device=dev, dtype=dt)
    out = buf[G:G + n_seq * N].unsqueeze(0)
    _lib.check(lib.avd_attention(qkv.data_ptr(), 0, out.data_ptr(), _lib.AVD_PREC_F16 if f16 else _lib.AVD_PREC_BF16,
                                 n_seq, N, H, hd, _lib.stream_ptr()))
    torch.cuda.synchronize()
    assert bool((buf[:G] == 7.0).all()) and bool((buf[G + n_seq * N:] == 7.0).all()), "attention wrote outside its output"
    # checker: fp32 softmax(q k^T / 8) v with torch on the GPU, one (sequence, head) at a time
    q, k, v = qkv.float().view(n_seq, N, 3, H, hd).permute(2, 0, 3, 1, 4)
    worst = 0.0
    got = out[0].float().view(n_seq, N, H, hd).permute(0, 2, 1, 3)
    for s in range(n_seq):
        for h in range(H):
            p = torch.softmax((q[s, h].double() @ k[s, h].double().t()) / math.sqrt(hd), dim=-1)
            worst = max(worst, rel_l2(got[s, h], p @ v[s, h].double()))
    print(f"[attention {'f16' if f16 else 'bf16'} n_seq={n_seq} N={N}] worst head rel-L2 {worst:.2e}")
    assert worst < tol


@pytest.mark.parametrize("n_seq,N", [(2, 133), (3, 445), (1, 192), (2, 129), (1, 2048)])
def test_attention_tcgen05_f16_template(dev, lib, n_seq, N):
    """The F16 instantiations of attn_tc_kernel (single-block N <= 192 and multi-block) at kernel level."""
    _attn_check(dev, lib, n_seq, N, True, 1.5e-3)


@pytest.mark.parametrize("f16", [True, False])
@pytest.mark.parametrize("N,ramp", [(1000, 10.0), (1000, 60.0), (2048, 25.0), (445, 30.0)])
def test_attention_multiblock_rising_max(dev, lib, N, ramp, f16):
    """The multi-block softmax takes its exponentials against the CARRIED row max and repeats a block only when the max
    rose by more than 2^8: scores that climb with the key index exercise both branches (ramp 10: ~2.7 log2 units per
    192-key block, repeats now and then; ramp 60: ~17 per block, repeats every block)."""
    from multimodal_diffusion_b200 import _lib
    with _lib.option("attn_fa", 0):           # attn_tc_kernel's two-tile multi-block path (the default is attn_fa_kernel)
        _attn_check(dev, lib, 2, N, f16, 1.5e-3 if f16 else 8e-3, key_ramp=ramp)


@pytest.mark.parametrize("n_seq,N,f16,ramp", [(3, 445, True, 0.0), (2, 193, False, 0.0), (1, 2048, True, 0.0),
                                              (2, 1000, False, 0.0), (1, 257, True, 0.0), (2, 384, False, 0.0),
                                              (1, 640, True, 0.0), (2, 1000, True, 10.0), (2, 1000, False, 60.0),
                                              (1, 2048, True, 25.0), (1, 16384, True, 0.0)])
@pytest.mark.parametrize("fa", [1, 2, 3, 4])
def test_attention_tmem_resident_kernel(dev, lib, n_seq, N, f16, ramp, fa):
    """attn_fa_kernel (option attn_fa; N > 192): P written to tensor memory and used as the A operand of P V, O accumulated
    in tensor memory and rescaled in place only when a row's max rises by more than 2^8, 128-key blocks through a 4-stage
    K / V ring.  Ragged last blocks (445 = 3 x 128 + 61, 193, 257: a one-row second query tile), scores that climb along
    the keys (in-place rescales in every block), 128 key blocks."""
    from multimodal_diffusion_b200 import _lib
    with _lib.option("attn_fa", fa):          # 1: two softmax warps per row group, 2: one thread per query row, 3: S double-buffered
        _attn_check(dev, lib, n_seq, N, f16, 1.5e-3 if f16 else 8e-3, key_ramp=ramp)


@pytest.mark.parametrize("fa", [1, 2, 3, 4])
def test_attention_tmem_resident_kernel_split_output(dev, lib, fa):
    """bf16x3 mode: bf16 q / k / v in, the attention output written as TWO bf16 segments (value + residual); the multi-block
    kernels' multi-segment epilogue against fp64, tighter than one bf16 segment allows."""
    from multimodal_diffusion_b200 import _lib
    H, hd, n_seq, N = 8, 64, 2, 445
    qkv = (randn(8, n_seq * N, 3 * H * hd) * 1.5).bfloat16().to(dev)
    out = torch.zeros(2, n_seq * N, H * hd, device=dev, dtype=torch.bfloat16)
    with _lib.option("attn_fa", fa):
        _lib.check(lib.avd_attention(qkv.data_ptr(), 0, out.data_ptr(), _lib.AVD_PREC_BF16X3, n_seq, N, H, hd,
                                     _lib.stream_ptr()))
    torch.cuda.synchronize()
    q, k, v = qkv.double().view(n_seq, N, 3, H, hd).permute(2, 0, 3, 1, 4)
    want = (torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(hd), dim=-1) @ v).permute(0, 2, 1, 3).reshape(n_seq * N, H * hd)
    one = rel_l2(out[0].float(), want)
    both = rel_l2(out.float().sum(0), want)
    print(f"[attention bf16x3 fa={fa}] rel-L2 one segment {one:.2e}, two segments {both:.2e}")
    assert both < 4e-3 and both < one                        # P is bf16: the second segment removes the output rounding only


@pytest.mark.parametrize("fa", [1, 2, 3, 4])
@pytest.mark.parametrize("N", [445, 1000])
def test_attention_tmem_resident_kernel_with_key_mask(dev, lib, N, fa):
    """key_padding_mask through attn_fa_kernel against the two-pass multi-block kernel: ragged tails, scattered padded keys,
    a fully padded 32-key chunk and a fully padded 128-key block."""
    from multimodal_diffusion_b200 import _lib, packing
    H, hd, n_seq = 8, 64, 3
    gen = torch.Generator().manual_seed(12)
    mask = torch.rand(n_seq, N, generator=gen) < 0.15
    for b in range(n_seq):
        mask[b, N - 1 - 9 * b:] = True
    mask[0, 32:64] = True
    mask[1, 128:256] = True
    mask[2, 0:130] = True                                # the first block has no key at all for this sequence
    bits = packing.pack_key_padding_mask(mask.to(dev))
    for dt, prec, tol in ((torch.float16, _lib.AVD_PREC_F16, 1.5e-3), (torch.bfloat16, _lib.AVD_PREC_BF16, 8e-3)):
        qkv = (randn(6, n_seq * N, 3 * H * hd) * 1.5).to(dt).to(dev)
        outs = []
        for opt in (0, fa):
            out = torch.zeros(1, n_seq * N, H * hd, device=dev, dtype=dt)
            with _lib.option("attn_fa", opt):
                _lib.check(lib.avd_attention_masked(qkv.data_ptr(), 0, out.data_ptr(), prec, n_seq, N, H, hd,
                                                    bits.data_ptr(), _lib.stream_ptr()))
            torch.cuda.synchronize()
            outs.append(out.float().cpu())
        assert torch.isfinite(outs[1]).all()
        assert rel_l2(outs[1], outs[0]) < tol


@pytest.mark.parametrize("n_seq,N,f16", [(3, 445, True), (2, 193, False), (1, 2048, True), (2, 1000, False)])
def test_attention_pipelined_kernel(dev, lib, n_seq, N, f16):
    """attn_pipe_kernel (option attn_pipe, off by default: measured slower than attn_tc_kernel's multi-block path, see
    its header): one query tile per CTA, S double-buffered in TMEM, S(j+2) and P V issued under the softmax of block j."""
    from multimodal_diffusion_b200 import _lib
    with _lib.option("attn_pipe", 1):
        _attn_check(dev, lib, n_seq, N, f16, 1.5e-3 if f16 else 8e-3)


@pytest.mark.parametrize("n_seq,N,f16", [(2, 133, True), (3, 133, False), (2, 136, True), (1, 128, False), (2, 100, True),
                                         (20, 133, True)])
def test_attention_small_kernel(dev, lib, n_seq, N, f16):
    """attn_small_kernel (two CTAs per SM, N <= 136: the rows past 128 run on CUDA cores), forced with attn_small = 2; the
    last case (160 (sequence, head) pairs > SM count) takes it by the default rule."""
    from multimodal_diffusion_b200 import _lib
    with _lib.option("attn_small", 2 if n_seq < 20 else 1):
        _attn_check(dev, lib, n_seq, N, f16, 1.5e-3 if f16 else 8e-3)


def test_attention_small_kernel_with_key_mask(dev, lib):
    """key_padding_mask through attn_small_kernel: the tile rows (TMEM path) and the thin rows (CUDA-core path) must ignore
    the same keys as the two-tile kernel."""
    from multimodal_diffusion_b200 import _lib, packing
    H, hd, n_seq, N = 8, 64, 3, 133
    qkv = (randn(5, n_seq * N, 3 * H * hd) * 1.5).to(torch.float16).to(dev)
    mask = torch.zeros(n_seq, N, dtype=torch.bool)
    mask[0, 100:] = True
    mask[1, 5:40] = True
    mask[2, ::3] = True
    bits = packing.pack_key_padding_mask(mask.to(dev))
    outs = []
    for opt in (0, 2):
        out = torch.zeros(1, n_seq * N, H * hd, device=dev, dtype=torch.float16)
        with _lib.option("attn_small", opt):
            _lib.check(lib.avd_attention_masked(qkv.data_ptr(), 0, out.data_ptr(), _lib.AVD_PREC_F16, n_seq, N, H, hd,
                                                bits.data_ptr(), _lib.stream_ptr()))
        torch.cuda.synchronize()
        outs.append(out.float().cpu())
    assert torch.isfinite(outs[1]).all()
    assert rel_l2(outs[1], outs[0]) < 1.5e-3


@pytest.mark.parametrize("f16", [True, False])
def test_attention_tcgen05_16k_tokens(dev, lib, f16):
    """configs[4]'s longest sequence: N = 16384 (86 key blocks)."""
    _attn_check(dev, lib, 1, 16384, f16, 2e-3 if f16 else 8e-3)


# ------------------------------------------------------------------------------------------ 1d: the drop-in entry points
class _FakeVAE:
    def __init__(self, z): self.z = z
    def encode(self, frames): return self.z
    def decode(self, z): return torch.zeros(1, 3, 1, 2, 2, device=z.device)


class _FakeCodec:
    def __init__(self, z): self.z = z
    def encode(self, wav): return self.z
    def decode(self, z): return torch.zeros(1, 1, 4, device=z.device)


@pytest.mark.parametrize("name,prompt_mod", [("v2a", "video"), ("a2v", "audio")])
def test_mirror_sample_one_direction_defaults_vs_reference(dev, mods, name, prompt_mod):
    """sample_clip.sample_one_direction / sample_latents of the mirror, DEFAULT arguments (f16/f16), stand-in encoders
    that inject the fixture's latents, against the reference's own run (tests/golden/sample_{v2a,a2v}.npz)."""
    from multimodal_diffusion_b200.avdiff.models.infer import sample_clip as sc
    cfg = json.loads((GOLDEN / f"cfg_mvp_{name}.json").read_text())
    g = golden(f"sample_{name}")
    zp, zi = T(g["z_prompt"]).to(dev), T(g["z_init"]).to(dev)
    seen = {}
    orig = sc.sample_latents

    def spy(**kw):
        seen["z"] = orig(**kw)
        return seen["z"]

    sc.sample_latents = spy
    try:
        out = sc.sample_one_direction(
            cfg=cfg, vid_vae=_FakeVAE(zp), aud_codec=_FakeCodec(zp), adapt_v=mods[0], adapt_a=mods[1], core=mods[2],
            head=mods[3], tstep_dim=256, prompt_modality=prompt_mod,
            prompt_video=np.zeros((48, 2, 2, 3), np.uint8) if prompt_mod == "video" else None,
            prompt_audio=np.zeros(16, np.float32) if prompt_mod == "audio" else None, device=dev, z_init=zi)
    finally:
        sc.sample_latents = orig
    assert ("audio" in out and out["sr"] == 16000) if prompt_mod == "video" else ("video" in out and out["fps"] == 16)
    err = rel_l2(seen["z"], g["z_final"])
    print(f"[mirror sample_one_direction {name}, defaults] final-latent rel-L2 {err:.3e}")
    assert err < TOL_FINAL
    # the FusedSampler behind it is cached: a second call with the same modules re-uses the captured graph
    n_cached = len(sc._SAMPLER_CACHE)
    z2 = sc.sample_latents(cfg=cfg, adapt_v=mods[0], adapt_a=mods[1], core=mods[2], head=mods[3],
                           target="audio" if prompt_mod == "video" else "video", z_prompt=zp, z_init=zi)
    assert len(sc._SAMPLER_CACHE) == n_cached and torch.equal(z2, seen["z"])


def test_mirror_build_components(dev):
    from multimodal_diffusion_b200.avdiff.models.infer import sample_clip as sc
    cfg = json.loads((GOLDEN / "cfg_mvp_v2a.json").read_text())
    vae, codec, adapt_v, adapt_a, core, head, tdim = sc.build_components(cfg, dev)
    assert type(vae).__name__ == "VideoVAE" and type(codec).__name__ == "AudioCodec" and tdim == 256
    assert codec.cfg.frames_per_clip == 150 and codec.hop == 320 and vae.cfg.t_down == 4 and vae.cfg.s_down == 8
    assert adapt_v.proj.weight.shape == (256, 256) and adapt_a.proj.weight.shape == (256, 32)
    assert len(core.blocks) == 8 and not core.training and not head.training
    ref = seeded_modules()
    for mine, theirs in zip((adapt_v, adapt_a, core, head), ref):          # reference state_dict keys load strictly
        mine.load_state_dict(theirs.state_dict(), strict=True)


# ------------------------------------------------------------------------------------------ 1f: fp16 range
def test_f16_range_check(dev, mods):
    """The fixed 2^-8 residual-stream scale covers |x| < 1.6e7; sample_checked scans every 16-bit operand buffer.
    Seeded latents pass; latents x10 either pass too or raise AvdRangeError -- never a silent wrong answer; values past
    the fp16 range must raise."""
    from multimodal_diffusion_b200.sampler import AvdRangeError
    B = 2
    fs = _sampler(mods, "audio", B, steps=6)
    zp, zi = randn(1, B, 8, 12, 16, 16).to(dev), randn(2, B, 8, 150).to(dev)
    z = fs.sample_checked(zp, zi)
    assert torch.equal(z, fs.sample_eager(zp, zi))                       # the scan does not change the result
    try:
        z10 = fs.sample_checked(zp * 10, zi * 10)
        assert torch.isfinite(z10).all()
        print("[f16 range] latents x10: no saturation")
    except AvdRangeError as e:
        print(f"[f16 range] latents x10: {e}")
    with pytest.raises(AvdRangeError):
        fs.sample_checked(zp * 1e9, zi * 1e9)
    # the scan kernel itself
    from multimodal_diffusion_b200 import _lib
    buf = torch.tensor([1.0, 65504.0, -65504.0, float("inf"), float("nan"), 3.0, -2.0, 60000.0], dtype=torch.float16,
                       device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    _lib.check(_lib.load().avd_f16_range_scan(buf.data_ptr(), buf.numel(), cnt.data_ptr(), _lib.stream_ptr()))
    assert int(cnt.item()) == 4


# ------------------------------------------------------------------------------------------ N4: eta > 0
def test_philox_normal_statistics_and_keys(dev, lib):
    from multimodal_diffusion_b200 import _lib
    n = 1 << 20
    out = torch.empty(4, n, device=dev)
    _lib.check(lib.avd_philox_normal(out.data_ptr(), 1234, 10, 7, 4, n, _lib.stream_ptr()))
    x = out.double()
    assert abs(float(x.mean())) < 3e-3 and abs(float(x.var()) - 1.0) < 5e-3
    assert abs(float((x ** 4).mean()) - 3.0) < 5e-2 and abs(float((x ** 3).mean())) < 2e-2
    assert float(x.abs().max()) < 7.0 and torch.isfinite(x).all()
    # a clip's noise depends on (seed, GLOBAL clip index, step, element) only
    one = torch.empty(1, n, device=dev)
    _lib.check(lib.avd_philox_normal(one.data_ptr(), 1234, 12, 7, 1, n, _lib.stream_ptr()))
    assert torch.equal(one[0], out[2])
    _lib.check(lib.avd_philox_normal(one.data_ptr(), 1234, 12, 8, 1, n, _lib.stream_ptr()))
    assert not torch.equal(one[0], out[2]) and abs(float((one[0].double() * out[2].double()).mean())) < 5e-3
    _lib.check(lib.avd_philox_normal(one.data_ptr(), 1235, 12, 7, 1, n, _lib.stream_ptr()))
    assert not torch.equal(one[0], out[2])


@pytest.mark.parametrize("target", ["audio", "video"])
def test_fused_loop_eta_positive(dev, mods, w_cpu, target):
    """ddim_eta > 0 inside the fused loop (schedule_utils.py:185-197): the step equals the reference formula with the
    device generator's z injected -- bit-exact given the GPU's own eps -- and the result is shard-invariant."""
    from multimodal_diffusion_b200 import _lib
    lib = _lib.load()
    B, eta, seed = 4, 0.7, 99
    fs = _sampler(mods, target, B, "fp32", "fp32", eta=eta, seed=seed, clip_index0=0)
    zp = randn(41, B, 8, 12, 16, 16) if target == "audio" else randn(41, B, 8, 150)
    zi = randn(42, B, 8, 150) if target == "audio" else randn(42, B, 8, 12, 16, 16)
    ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5, eta=eta)
    ab, ts = ocfg.alpha_bar(), ocfg.timesteps()
    fs.prepare(zp.to(dev))
    z = zi.to(dev).clone()
    tap = torch.empty(2 * B, fs.Nt, fs.Dt, device=dev)
    per_clip = zi[0].numel()
    for i in range(3):
        z_t = z.cpu().clone()
        fs.step(i, z, tap)
        noise = torch.empty(B, per_clip, device=dev)
        _lib.check(lib.avd_philox_normal(noise.data_ptr(), seed, 0, i, B, per_clip, _lib.stream_ptr()))
        ec, en = tap[:B].cpu(), tap[B:].cpu()
        e_hat = en + 3.5 * (ec - en)
        want = orc.ddim_step(z_t, ts[i].repeat(B), ts[i + 1].repeat(B), orc.untokenise(ocfg, target, e_hat, z_t), ab,
                             eta=eta, noise=noise.cpu().view_as(z_t))
        assert torch.equal(z.cpu(), want), f"step {i}: fused eta>0 update is not the reference formula"
    # per-step eps still matches the oracle (the noise does not enter eps)
    with torch.no_grad():
        z_v, z_a = (zp, zi) if target == "audio" else (zi, zp)
        e_c, e_n, _ = orc.eps_for_step(ocfg, w_cpu, target, z_v, z_a, ts[0].repeat(B))
    z0 = zi.to(dev).clone()
    fs.step(0, z0, tap)
    assert rel_l2(tap[:B], e_c) < TOL_EPS_FP32 and rel_l2(tap[B:], e_n) < TOL_EPS_FP32
    # shard invariance: clips 2,3 sampled alone with clip_index0 = 2 give the same latents as inside the batch of 4
    full = _sampler(mods, target, B, "fp32", "fp32", steps=4, eta=eta, seed=seed).sample_eager(zp.to(dev), zi.to(dev))
    part = _sampler(mods, target, 2, "fp32", "fp32", steps=4, eta=eta, seed=seed, clip_index0=2) \
        .sample_eager(zp[2:].to(dev), zi[2:].to(dev))
    err_same = rel_l2(part, full[2:])
    assert err_same < 1e-5
    other = _sampler(mods, target, 2, "fp32", "fp32", steps=4, eta=eta, seed=seed, clip_index0=0) \
        .sample_eager(zp[2:].to(dev), zi[2:].to(dev))
    # a different clip index draws different noise (sigma * z is small against |z| ~ 1e4 after 4 steps, but far above
    # the fp32 agreement of the same-index run)
    assert rel_l2(other, full[2:]) > max(2e-5, 5 * err_same)


# ------------------------------------------------------------------------------------------ N4: layernorm core, head variants
@pytest.mark.parametrize("core_p,head_p,tol", [("fp32", "fp32", TOL_EPS_FP32), ("f16", "f16", TOL_EPS_16)])
def test_fused_sampler_layernorm_core(dev, mods, core_p, head_p, tol):
    """MMDiT(norm="layernorm") (mmdt.py:45-46) inside the fused sampler: LayerNorm passes in front of QKV / fc1 and the
    final nn.LayerNorm in front of the head."""
    from multimodal_diffusion_b200.avdiff.models.mmdt import MMDiT
    from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig
    torch.manual_seed(5)
    core = MMDiT(d_model=512, n_layers=3, n_heads=8, mlp_ratio=4.0, dropout=0.0, norm="layernorm").eval()
    with torch.no_grad():
        for k, v in core.state_dict().items():
            if "norm" in k:
                v.copy_(1.0 + 0.2 * randn(len(k), *v.shape) if k.endswith("weight") else 0.1 * randn(len(k) + 1, *v.shape))
    w = weights_dict()
    w["core"] = {k: v.clone() for k, v in core.state_dict().items()}
    B = 2
    cfg = HotPathConfig(sampler_steps=50, guidance=3.5, n_layers=3)
    fs = FusedSampler(cfg, mods[0], mods[1], core.to(dev), mods[3], "audio", B, device=dev, core_precision=core_p,
                      head_precision=head_p)
    ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5, n_layers=3)
    zp, zi = randn(41, B, 8, 12, 16, 16), randn(42, B, 8, 150)
    fs.prepare(zp.to(dev))
    ec, en, ehat, ez = _step_vs_oracle(fs, ocfg, w, "audio", zi, zp, 0, dev)
    print(f"[layernorm core {core_p}] eps cond/null {ec:.2e}/{en:.2e} cfg {ehat:.2e}")
    assert max(ec, en) < tol
    z_g, z_e = fs.sample(zp.to(dev), zi.to(dev)), fs.sample_eager(zp.to(dev), zi.to(dev))
    assert torch.equal(z_g, z_e) and torch.isfinite(z_g).all()


@pytest.mark.parametrize("name", ["leaky_spec2", "relu_shared_spec", "gelu_noshared"])
def test_head_variants_vs_reference(dev, name):
    """MultiModalNoiseHead with modality-specific trunks / share_parameters / relu / leaky_relu (noise_heads.py:110-229)
    against outputs of the reference's own module."""
    from test_round2_host import make_head_variant
    g = golden("head_variants")
    head, _ = make_head_variant(name)
    head = head.to(dev)
    head.precision = "fp32"
    h = T(g["h"]).to(dev)
    with torch.no_grad():
        out = head({"video": h, "audio": h})
    for m in ("video", "audio"):
        assert rel_l2(out[m], g[f"{name}_{m}"]) < TOL_EPS_FP32, (name, m)


def test_fused_sampler_head_with_specific_trunk(dev, mods, w_cpu):
    """A head with a leaky_relu activation and two modality-specific blocks through the fused loop (fp32 mode)."""
    from multimodal_diffusion_b200.avdiff.models.heads.noise_heads import MultiModalNoiseHead
    from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig
    torch.manual_seed(77)
    head = MultiModalNoiseHead(input_dims={"video": 512, "audio": 512}, output_dims={"video": 256, "audio": 32},
                               hidden_dim=512, num_shared_layers=1, num_modality_specific_layers=3, dropout=0.1,
                               activation="leaky_relu").eval()
    w = dict(w_cpu, head={k: v.clone() for k, v in head.state_dict().items()})
    B = 2
    fs = FusedSampler(HotPathConfig(sampler_steps=50, guidance=3.5), mods[0], mods[1], mods[2], head.to(dev), "audio", B,
                      device=dev, core_precision="fp32", head_precision="fp32")
    assert fs.head.struct.n_shared == 3 and fs.head.struct.activation == 2
    ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5, head_activation="leaky_relu")
    zp, zi = randn(41, B, 8, 12, 16, 16), randn(42, B, 8, 150)
    fs.prepare(zp.to(dev))
    ec, en, ehat, ez = _step_vs_oracle(fs, ocfg, w, "audio", zi, zp, 0, dev)
    assert max(ec, en) < TOL_EPS_FP32
    # a training-mode head with dropout is refused, a bare state_dict needs the activation spelled out
    with pytest.raises(NotImplementedError, match="training mode"):
        FusedSampler(HotPathConfig(), mods[0], mods[1], mods[2], head.train(), "audio", B, device=dev)
    head.eval()
    with pytest.raises(ValueError, match="activation"):
        FusedSampler(HotPathConfig(), mods[0], mods[1], mods[2], head.state_dict(), "audio", B, device=dev)


# ------------------------------------------------------------------------------------------ N3 / N4: trainer path, additive tables
def test_fused_sampler_trainer_checkpoint_additive_time(dev, mods, w_cpu):
    """N3: a trainer-format checkpoint (full-width adapters, ADDED timestep embedding, EMA shadow) sampled through
    checkpoint.sampler_from_checkpoint; first steps against the oracle's additive path (pinned by trainer_path.npz)."""
    from multimodal_diffusion_b200 import checkpoint as ck
    from multimodal_diffusion_b200.avdiff.models.infer.sample_clip import LinearAdapter
    from multimodal_diffusion_b200.sampler import HotPathConfig
    torch.manual_seed(400); av = LinearAdapter(256, 512)
    torch.manual_seed(401); aa = LinearAdapter(32, 512)
    core_sd = w_cpu["core"]
    ema = {"module." + k: v * 1.01 for k, v in core_sd.items()}          # a shadow that differs from the raw weights
    state = {"step": 10, "core": core_sd, "head": w_cpu["head"], "adapt_v": av.state_dict(), "adapt_a": aa.state_dict(),
             "vid_vae": {}, "aud_codec": {}, "opt": {}, "ema": ema}
    B = 2
    fs = ck.sampler_from_checkpoint(HotPathConfig(sampler_steps=50, guidance=3.5), state, "audio", B, device=dev,
                                    core_precision="fp32", head_precision="fp32")
    w = {"adapt_v": av.state_dict(), "adapt_a": aa.state_dict(), "head": w_cpu["head"],
         "core": {k: v * 1.01 for k, v in core_sd.items()}}
    ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5, time_embedding="add")
    zp, zi = randn(41, B, 8, 12, 16, 16), randn(42, B, 8, 150)
    fs.prepare(zp.to(dev))
    z_or = zi.clone()
    ab, ts = ocfg.alpha_bar(), ocfg.timesteps()
    for i in range(2):
        ec, en, ehat, ez = _step_vs_oracle(fs, ocfg, w, "audio", z_or, zp, i, dev)
        print(f"[trainer ckpt, additive time, step {i}] eps cond/null {ec:.2e}/{en:.2e} z_prev {ez:.2e}")
        assert max(ec, en) < TOL_EPS_FP32
        with torch.no_grad():
            _, _, e_hat = orc.eps_for_step(ocfg, w, "audio", zp, z_or, ts[i].repeat(B))
            z_or = orc.ddim_step(z_or, ts[i].repeat(B), ts[i + 1].repeat(B), orc.untokenise(ocfg, "audio", e_hat, z_or), ab)


@pytest.mark.parametrize("target", ["audio", "video"])
def test_fused_sampler_positional_and_modality_embeddings(dev, mods, w_cpu, target):
    """N4(iii): ModalityEmbedding + PositionalEmbedding3D/1D rows added in the adapter epilogue of the fused input
    builder (FusedSampler(add_v, add_a)) against the oracle applying the same tables after the adapters."""
    from multimodal_diffusion_b200.avdiff.models import adapters as ad
    torch.manual_seed(200); me = ad.ModalityEmbedding(256)
    torch.manual_seed(202); p3 = ad.PositionalEmbedding3D(256, max_T=8, max_H=8, max_W=8, mode="learned")
    p1 = ad.PositionalEmbedding1D(256, mode="sin")
    add_v = ad.token_additive_table(modality_embedding=me, modality="video", positional=p3, grid=(6, 4, 4))
    add_a = ad.token_additive_table(modality_embedding=me, modality="audio", positional=p1, grid=(37,))
    B = 2
    fs = _sampler(mods, target, B, "fp32", "fp32", add_v=add_v, add_a=add_a)
    ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5, add_v=add_v, add_a=add_a)
    zp = randn(41, B, 8, 12, 16, 16) if target == "audio" else randn(41, B, 8, 150)
    zi = randn(42, B, 8, 150) if target == "audio" else randn(42, B, 8, 12, 16, 16)
    fs.prepare(zp.to(dev))
    ec, en, ehat, ez = _step_vs_oracle(fs, ocfg, w_cpu, target, zi, zp, 0, dev)
    assert max(ec, en) < TOL_EPS_FP32
    plain = _sampler(mods, target, B, "fp32", "fp32")
    plain.prepare(zp.to(dev))
    ec0, _, _, _ = _step_vs_oracle(plain, ocfg, w_cpu, target, zi, zp, 0, dev)
    assert ec0 > 100 * TOL_EPS_FP32                                        # the tables matter


# ------------------------------------------------------------------------------------------ schedule variants of the same path
def test_build_x_wide_kernel_bit_identical(dev, mods):
    """The 32-token build_x kernel (A2V video rows at batch >= 4) sums k in the same order as the 4-token one."""
    from multimodal_diffusion_b200 import _lib
    B = 8
    zp, zi = randn(61, B, 8, 150).to(dev), randn(62, B, 8, 12, 16, 16).to(dev)
    outs = []
    for wide in (0, 1):
        with _lib.option("bx_wide", wide):
            fs = _sampler(mods, "video", B, "fp32", "fp32", steps=2)
            outs.append(fs.sample_eager(zp, zi))
    assert torch.equal(outs[0], outs[1])


@pytest.mark.parametrize("core_precision", ["fp32", "f16"])
def test_build_x_tensor_core_route(dev, mods, w_cpu, core_precision):
    """Option bx_wide = 2: tokenise -> tcgen05 GEMM (split-bf16, ~fp32 products) -> assemble.  Against the FMA kernel on
    the same inputs (2-step loop) and, in fp32, against the oracle's first step at the fp32 tolerance."""
    from multimodal_diffusion_b200 import _lib
    B = 8
    zp, zi = randn(61, B, 8, 150), randn(62, B, 8, 12, 16, 16)
    outs = []
    for opt in (0, 2):
        with _lib.option("bx_wide", opt):
            fs = _sampler(mods, "video", B, core_precision, core_precision, steps=2)
            outs.append(fs.sample_eager(zp.to(dev), zi.to(dev)))
    err = rel_l2(outs[1], outs[0])
    print(f"build_x tc vs fma ({core_precision}): {err:.2e}")
    # a 2-step schedule multiplies eps differences by ~1/sqrt(alpha_bar_999); in f16 the ~1e-7 input difference flips
    # fp16 roundings downstream, so the f16 case is held to the oracle tolerance instead
    if core_precision == "fp32":
        assert err < 2e-6
    with _lib.option("bx_wide", 2):
        fs = _sampler(mods, "video", B, core_precision, core_precision)
        ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5)
        fs.prepare(zp.to(dev))
        ec, en, ehat, ez = _step_vs_oracle(fs, ocfg, w_cpu, "video", zi, zp, 0, dev)
    print(f"build_x tc vs oracle ({core_precision}): eps {max(ec, en):.2e} eps_hat {ehat:.2e}")
    assert max(ec, en) < (TOL_EPS_FP32 if core_precision == "fp32" else TOL_EPS_16)


def test_chains_schedule_matches_single_stream(dev, mods, w_cpu):
    """Option chains = 2: the sequence groups run on forked streams (eager and inside a captured graph); same numbers
    as the single-stream schedule up to the tile shapes the narrower GEMMs pick, and within tolerance of the oracle."""
    from multimodal_diffusion_b200 import _lib
    B = 8
    zp, zi = randn(1, B, 8, 12, 16, 16), randn(2, B, 8, 150)
    base = _sampler(mods, "audio", B, steps=4).sample(zp.to(dev), zi.to(dev))
    with _lib.option("chains", 2):
        fs = _sampler(mods, "audio", B, steps=4)
        z_graph = fs.sample(zp.to(dev), zi.to(dev))
        z_eager = fs.sample_eager(zp.to(dev), zi.to(dev))
        fs50 = _sampler(mods, "audio", B)
        ocfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5)
        fs50.prepare(zp.to(dev))
        ec, en, ehat, ez = _step_vs_oracle(fs50, ocfg, w_cpu, "audio", zi, zp, 0, dev)
    torch.cuda.synchronize()
    assert torch.equal(z_graph, z_eager)
    assert rel_l2(z_graph, base) < 5e-3
    assert max(ec, en) < TOL_EPS_16 and ehat < TOL_EPS_16


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_fused_sampler_on_non_current_device(mods):
    """FusedSampler(device="cuda:1") while the current device is cuda:0 launches everything on cuda:1."""
    B = 2
    zp, zi = randn(1, B, 8, 12, 16, 16), randn(2, B, 8, 150)
    z0 = _sampler(mods, "audio", B, steps=3).sample(zp.to("cuda:0"), zi.to("cuda:0"))
    from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig
    assert torch.cuda.current_device() == 0
    fs1 = FusedSampler(HotPathConfig(sampler_steps=3, guidance=3.5), *[m for m in seeded_modules()], "audio", B,
                       device="cuda:1")
    z1 = fs1.sample(zp.to("cuda:1"), zi.to("cuda:1"))
    z1e = fs1.sample_eager(zp.to("cuda:1"), zi.to("cuda:1"))
    assert z1.device.index == 1 and torch.equal(z1, z1e) and torch.equal(z1.cpu(), z0.cpu())
    with pytest.raises(ValueError, match="lives on"):
        fs1.sample_eager(zp.to("cuda:0"), zi.to("cuda:0"))


# ------------------------------------------------------------------------------------------ N1: encoders / decoders
def _golden_codec(seed, **kw):
    from multimodal_diffusion_b200.avdiff.models.encoders.audio_codec import AudioCodec, AudioCodecConfig
    torch.manual_seed(seed)
    return AudioCodec(AudioCodecConfig(lat_ch=8, hidden=64, smooth_kernel=7, **kw)).eval()


def _golden_vae():
    from multimodal_diffusion_b200.avdiff.models.encoders.vae_video3d import VideoVAE, VideoVAEConfig
    torch.manual_seed(502)
    vae = VideoVAE(VideoVAEConfig(lat_ch=8, t_down=4, s_down=8, enc_base=64, enc_blocks=2, dec_base=64, dec_blocks=2)).eval()
    with torch.no_grad():
        for k, v in vae.state_dict().items():
            if ".2.weight" in k:
                v.copy_(1.0 + 0.2 * randn(len(k), *v.shape))
            elif ".2.bias" in k:
                v.copy_(0.1 * randn(len(k) + 1, *v.shape))
    return vae


def test_audio_codec_vs_reference(dev):
    """AudioCodec.encode / decode (audio_codec.py:184-214) against the reference module's outputs: exact-frame pooling
    with a right pad (L = 3163 -> 10 frames of 317), the hop_samples path (ceil(L / hop) frames), nearest upsample."""
    g = golden("encoders")
    codec = _golden_codec(500, hop_samples=320, frames_per_clip=10)
    sums = np.array([[float(v.double().sum()), float(v.double().abs().sum())] for v in codec.state_dict().values()])
    assert np.allclose(sums, g["codec_sums"], rtol=1e-12)
    codec = codec.to(dev)
    with torch.no_grad():
        z = codec.encode(T(g["wav"]).to(dev))
        assert z.shape == (2, 8, 10) and rel_l2(z, g["z_a"]) < TOL_EPS_FP32
        wav_hat = codec.decode(T(g["z_a"]).to(dev))
        assert wav_hat.shape == (2, 1, 3200) and rel_l2(wav_hat, g["wav_hat"]) < TOL_EPS_FP32
        codec2 = _golden_codec(501, hop_samples=160, frames_per_clip=None).to(dev)
        z2 = codec2.encode(T(g["wav2"]).to(dev))
        assert z2.shape == (1, 8, 11) and rel_l2(z2, g["z_a2"]) < TOL_EPS_FP32
        assert rel_l2(codec2.decode(T(g["z_a2"]).to(dev)), g["wav_hat2"]) < TOL_EPS_FP32
        with pytest.raises(AssertionError):
            codec.encode(torch.zeros(2, 2, 100, device=dev))                  # audio_codec.py:191
    from multimodal_diffusion_b200 import _lib
    with pytest.raises(_lib.AvdError, match="CUDA"), torch.no_grad():
        codec.encode(torch.zeros(1, 1, 640))                                  # no CPU path


def test_video_vae_vs_reference(dev):
    """VideoVAE.encode / decode (vae_video3d.py:164-214) against the reference module: conv3d + GELU + GroupNorm blocks,
    AvgPool3d, trilinear upsample; plus the center-crop of a non-divisible clip and a round trip at mvp size."""
    g = golden("encoders")
    vae = _golden_vae()
    sums = np.array([[float(v.double().sum()), float(v.double().abs().sum())] for v in vae.state_dict().values()])
    assert np.allclose(sums, g["vae_sums"], rtol=1e-12)
    vae = vae.to(dev)
    x = T(g["frames"]).to(dev)
    with torch.no_grad():
        z = vae.encode(x)
        err_e = rel_l2(z, g["z_v"])
        x_hat = vae.decode(T(g["z_v"]).to(dev))
        err_d = rel_l2(x_hat, g["frames_hat"])
        print(f"[VideoVAE] encode rel-L2 {err_e:.2e}  decode rel-L2 {err_d:.2e}")
        assert z.shape == (1, 8, 2, 4, 4) and err_e < TOL_EPS_FP32
        assert x_hat.shape == (1, 3, 8, 32, 32) and err_d < TOL_EPS_FP32
        # non-divisible input: center-cropped like the reference (33 -> 32 rows from offset 0, 9 -> 8 frames)
        xp = torch.nn.functional.pad(x, (0, 0, 0, 1, 0, 1))
        with pytest.warns(UserWarning):
            zc = vae.encode(xp)
        assert torch.equal(zc, z)
        # full mvp size (3 s, 128 x 128): shapes, finiteness, value range, determinism
        big = torch.rand(1, 3, 48, 128, 128, device=dev)
        zb = vae.encode(big)
        xb = vae.decode(zb)
        assert zb.shape == (1, 8, 12, 16, 16) and xb.shape == (1, 3, 48, 128, 128)
        assert torch.isfinite(xb).all() and float(xb.min()) >= 0.0 and float(xb.max()) <= 1.0
        assert torch.equal(vae.encode(big), zb)


def test_sample_one_direction_end_to_end_with_encoders(dev, mods):
    """wav / frames in -> frames / wav out with NO caller-supplied encoder: build_components' own VideoVAE / AudioCodec
    (CUDA kernels), the fused loop, the decoders (sample_clip.py:220-394)."""
    from multimodal_diffusion_b200.avdiff.models.infer import sample_clip as sc
    cfg = json.loads((GOLDEN / "cfg_mvp_v2a.json").read_text())
    cfg["diffusion"]["audio"]["sampler_steps"] = 4
    cfg["diffusion"]["video"]["sampler_steps"] = 4
    torch.manual_seed(9)
    vae, codec, _, _, _, _, tdim = sc.build_components(cfg, dev)
    frames = (torch.rand(48, 128, 128, 3) * 255).to(torch.uint8).numpy()
    out = sc.sample_one_direction(cfg=cfg, vid_vae=vae, aud_codec=codec, adapt_v=mods[0], adapt_a=mods[1], core=mods[2],
                                  head=mods[3], tstep_dim=tdim, prompt_modality="video", prompt_video=frames,
                                  prompt_audio=None, device=dev)
    assert out["audio"].shape == (48000,) and out["sr"] == 16000 and np.isfinite(out["audio"]).all()
    assert np.abs(out["audio"]).max() <= 1.0
    wav = np.tanh(np.random.default_rng(3).standard_normal(48000)).astype(np.float32)
    out = sc.sample_one_direction(cfg=cfg, vid_vae=vae, aud_codec=codec, adapt_v=mods[0], adapt_a=mods[1], core=mods[2],
                                  head=mods[3], tstep_dim=tdim, prompt_modality="audio", prompt_video=None,
                                  prompt_audio=wav, device=dev)
    assert out["video"].shape == (48, 128, 128, 3) and out["video"].dtype == np.uint8 and out["fps"] == 16


# ------------------------------------------------------------------------------------------ chained GEMM launch
def test_chained_block_launch_matches_unchained(dev, lib):
    """avd_gemm_chain.cu (option gemm_chain, off by default): out-proj -> fc1 -> fc2 -> next QKV of every block as ONE
    persistent launch with row-group flags.  Same tiles and the same per-tile arithmetic as the four separate launches
    except fc1's accumulation order (one accumulator in k order instead of two alternating ones), so the results agree to
    fp16 rounding flips; it is deterministic, graph replay equals eager, and it leaves its flag block zero."""
    from multimodal_diffusion_b200 import _lib
    from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig
    mods = seeded_modules()
    B = 8
    outs, launches = {}, {}
    for chain in (0, 1, 1):
        with _lib.option("gemm_chain", chain):
            fs = FusedSampler(HotPathConfig(sampler_steps=4, guidance=3.5), mods[0], mods[1], mods[2], mods[3], "audio", B,
                              device=dev)
            zp, z0 = randn(1, *fs.prompt_shape).to(dev), randn(2, *fs.latent_shape).to(dev)
            n0 = lib.avd_launch_count()
            ze = fs.sample_eager(zp, z0)
            launches[chain] = lib.avd_launch_count() - n0
            zg = fs.sample(zp, z0)
            torch.cuda.synchronize()
            fs.check_errors()
            assert torch.equal(ze, zg)
            outs.setdefault(chain, []).append(ze.clone())
    # per step: blocks 0..6 fold {out-proj, fc1, fc2, next QKV} into one launch, the last block {out-proj, fc1, fc2}
    assert launches[0] - launches[1] == 4 * (7 * 3 + 2)
    assert torch.equal(outs[1][0], outs[1][1])
    assert rel_l2(outs[1][0].cpu(), outs[0][0].cpu()) < 5e-3
