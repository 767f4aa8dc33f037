"""GPU: each kernel of libavdiff_b200.so through the C ABI / the mirror ops, against the CPU oracle
(bit-exact for data movement and the fp32 elementwise chains; stated tolerances for the GEMM / attention)."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

from _common import golden, orc, randn, rel_l2

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev(lib):
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from multimodal_diffusion_b200 import _lib
    _lib.check(lib.avd_device_check())
    return torch.device("cuda:0")


def T(a):
    return torch.from_numpy(np.asarray(a))


# ------------------------------------------------------------------------------------------ tokenisers
@pytest.mark.parametrize("shape,tube", [((2, 8, 12, 16, 16), (2, 4, 4)), ((1, 3, 4, 6, 8), (2, 3, 2)),
                                        ((1, 3, 4, 6, 8), (1, 2, 1)), ((3, 8, 40, 16, 16), (2, 4, 4)),
                                        ((0, 8, 12, 16, 16), (2, 4, 4))])
def test_tube_patch_unpatch_bit_exact(dev, shape, tube):
    from multimodal_diffusion_b200.avdiff.utils import ops
    z = randn(1, *shape)
    t, h, w = tube
    want = orc.tube_patch_video(z, t, h, w)
    got = ops.tube_patch_video(z.to(dev), t, h, w)
    assert got.shape == want.shape and torch.equal(got.cpu(), want)
    back = ops.tube_unpatch_video(got, shape[1], shape[2], shape[3], shape[4], t, h, w)
    assert torch.equal(back.cpu(), z)                      # the reference's own round-trip pin (test_shapes.py:26-36)


def test_tube_golden_and_errors(dev):
    from multimodal_diffusion_b200.avdiff.utils import ops
    g = golden("tokenizers")
    zs = T(g["tube_small_z"]).to(dev)
    assert np.array_equal(ops.tube_patch_video(zs, 2, 3, 2).cpu().numpy(), g["tube_small_tok_t2h3w2"])
    assert np.array_equal(ops.tube_patch_video(zs, 1, 2, 1).cpu().numpy(), g["tube_small_tok_t1h2w1"])
    with pytest.raises(AssertionError):
        ops.tube_patch_video(zs, 3, 3, 2)                  # ops.py:113
    with pytest.raises(AssertionError):
        ops.tube_unpatch_video(torch.zeros(1, 5, 7, device=dev), 3, 4, 6, 8, 2, 3, 2)   # ops.py:137


def test_audio_chunk_unchunk_bit_exact(dev):
    from multimodal_diffusion_b200.avdiff.utils import ops
    from multimodal_diffusion_b200.avdiff.models.tokenizers import AudioTokenizer, AudioTokenizerCfg
    g = golden("tokenizers")
    za = T(g["aud_z"]).to(dev)
    assert np.array_equal(ops.chunk_1d(za, 4, 4).cpu().numpy(), g["aud_windows_l4s4"])
    assert np.array_equal(ops.chunk_1d(za, 6, 2).cpu().numpy(), g["aud_windows_l6s2"])
    assert np.array_equal(ops.latents_to_tokens_audio(za, 4, 4).cpu().numpy(), g["aud_tok_l4s4"])
    assert np.array_equal(ops.latents_to_tokens_audio(za, 6, 2).cpu().numpy(), g["aud_tok_l6s2"])
    assert np.array_equal(AudioTokenizer().encode(za).cpu().numpy(), g["aud_tok_l4s4"])
    tk = T(g["aud_untok_in"]).to(dev)
    u = ops.tokens_to_latents_audio(tk, 8, 4, 150, 4)
    assert np.array_equal(u.cpu().numpy(), g["aud_untok_l4s4_F150"])
    assert float(u[..., 148:].abs().max()) == 0.0
    assert np.array_equal(ops.tokens_to_latents_audio(tk, 8, 4, 140, 4).cpu().numpy(), g["aud_untok_l4s4_F140"])
    tk2 = T(g["aud_untok2_in"]).to(dev)
    assert np.array_equal(AudioTokenizer(AudioTokenizerCfg(6, 2)).decode(tk2, 8, 150).cpu().numpy(),
                          g["aud_untok_l6s2_F150"])
    w = T(g["ola_in"]).to(dev)
    for s in (3, 5, 7):
        assert np.array_equal(ops.overlap_add_1d(w, stride=s).cpu().numpy(), g[f"ola_s{s}"])
    short = ops.chunk_1d(randn(6, 2, 3).to(dev), 4, 4)
    assert np.array_equal(short.cpu().numpy(), g["chunk_short"])
    # transposed dim + long clip (10 s config: Fa = 500)
    x = randn(9, 2, 500, 8)
    assert torch.equal(ops.chunk_1d(x.to(dev), 4, 4, dim=1).cpu(), orc.chunk_1d(x, 4, 4, dim=1))


# ------------------------------------------------------------------------------------------ schedules
def test_ddim_step_bit_exact(dev):
    from multimodal_diffusion_b200.avdiff.utils import schedule_utils as su
    from multimodal_diffusion_b200.avdiff.models.schedules import ModalitySchedule
    g = golden("schedules")
    ab = T(g["alpha_bar_cosine"])
    assert np.array_equal(su.alphas_cumprod_from_betas(su.make_beta_schedule(1000, "cosine"))[1].numpy(),
                          g["alpha_bar_cosine"])
    x, e = T(g["ddim_x"]).to(dev), T(g["ddim_eps"]).to(dev)
    for name, (tn, tp) in {"first": (999, 979), "mid": (499, 479), "last": (19, -1)}.items():
        out = su.ddim_step(x, torch.full((3,), tn), torch.full((3,), tp), e, ab)
        assert np.array_equal(out.cpu().numpy(), g[f"ddim_{name}"]), name
    out = su.ddim_step(x, torch.tensor([999, 499, 19]), torch.tensor([979, 479, -1]), e, ab.to(dev))
    assert np.array_equal(out.cpu().numpy(), g["ddim_mixed_t"])
    sch = ModalitySchedule.make(kind="cosine", steps=1000).to(dev)
    out2 = sch.ddim_step(x, torch.tensor([999, 499, 19], device=dev), torch.tensor([979, 479, -1], device=dev), e)
    assert np.array_equal(out2.cpu().numpy(), g["ddim_mixed_t"])
    # eta > 0: same formula with an injected noise tensor (statistical check of the sigma * z term)
    xs, es = randn(3, 4, 8, 150).to(dev), randn(4, 4, 8, 150).to(dev)
    o = su.ddim_step(xs, torch.full((4,), 499), torch.full((4,), 479), es, ab, eta=1.0)
    base = orc.ddim_step(xs.cpu(), torch.full((4,), 499), torch.full((4,), 479), es.cpu(), ab, eta=1.0,
                         noise=torch.zeros(4, 8, 150))
    a_t, a_p = float(ab[499]), float(ab[479])
    sigma = math.sqrt((1 - a_p) / (1 - a_t) * (1 - a_t / a_p))
    resid = (o.cpu() - base) / sigma
    assert abs(float(resid.mean())) < 0.06 and abs(float(resid.std()) - 1.0) < 0.06


def test_timestep_embedding(dev):
    from multimodal_diffusion_b200.avdiff.utils import schedule_utils as su
    g = golden("schedules")
    ts = su.make_sampling_schedule(1000, 50)
    assert np.array_equal(ts.numpy(), g["sched_50"])
    assert np.array_equal(su.timestep_embedding_host(ts[:-1], 256).numpy(), g["temb_steps_256"])   # host table: exact
    emb = su.timestep_embedding(ts[:-1].to(dev), 256).cpu().numpy()
    assert np.abs(emb - g["temb_steps_256"]).max() < 2e-4       # device sin/cos of arguments up to 999 rad
    odd = su.timestep_embedding(torch.tensor([0, 3, 999], device=dev), 7).cpu().numpy()
    assert odd.shape == (3, 7) and np.abs(odd - g["temb_odd_7"]).max() < 2e-4


# ------------------------------------------------------------------------------------------ fused CFG + untokenise + DDIM
@pytest.mark.parametrize("B", [1, 8])
@pytest.mark.parametrize("tn,tp", [(999, 979), (19, -1)])
def test_cfg_untok_ddim_video_bit_exact(dev, lib, B, tn, tp):
    from multimodal_diffusion_b200 import _lib
    cfg = orc.HotPathConfig()
    ab = cfg.alpha_bar()
    ec, en, z = randn(1, B, 96, 256) * 50, randn(2, B, 96, 256) * 50, randn(3, B, 8, 12, 16, 16) * 30
    e_hat = en + 3.5 * (ec - en)
    want = orc.ddim_step(z, torch.full((B,), tn), torch.full((B,), tp),
                         orc.tube_unpatch_video(e_hat, 8, 12, 16, 16, 2, 4, 4), ab)
    a = _lib.CfgDdimArgs(target=_lib.AVD_TARGET_VIDEO, B=B, C=8, T=12, H=16, W=16, tt=2, th=4, tw=4, guidance=3.5,
                         t_now=tn, t_prev=tp, T_train=1000)
    zd, out = z.to(dev), torch.empty(B, 8, 12, 16, 16, device=dev)
    ecd, end, abd = ec.to(dev), en.to(dev), ab.to(dev)      # keep the device tensors alive across the launch
    _lib.check(lib.avd_cfg_untok_ddim(C.byref(a), ecd.data_ptr(), end.data_ptr(), zd.data_ptr(), out.data_ptr(),
                                      abd.data_ptr(), _lib.stream_ptr()))
    assert torch.equal(out.cpu(), want)


@pytest.mark.parametrize("length,stride,Fa", [(4, 4, 150), (6, 2, 150), (4, 4, 500)])
def test_cfg_untok_ddim_audio_bit_exact(dev, lib, length, stride, Fa):
    from multimodal_diffusion_b200 import _lib
    B, Ca = 3, 8
    Na = (Fa - length) // stride + 1
    ab = orc.HotPathConfig().alpha_bar()
    ec, en, z = randn(1, B, Na, Ca * length), randn(2, B, Na, Ca * length), randn(3, B, Ca, Fa)
    e_hat = en + 3.5 * (ec - en)
    want = orc.ddim_step(z, torch.full((B,), 979), torch.full((B,), 959),
                         orc.tokens_to_latents_audio(e_hat, Ca, length, Fa, stride), ab)
    a = _lib.CfgDdimArgs(target=_lib.AVD_TARGET_AUDIO, B=B, C=Ca, L=Fa, length=length, stride=stride, Nt=Na,
                         guidance=3.5, t_now=979, t_prev=959, T_train=1000)
    zd, ecd, end, abd = z.to(dev), ec.to(dev), en.to(dev), ab.to(dev)
    _lib.check(lib.avd_cfg_untok_ddim(C.byref(a), ecd.data_ptr(), end.data_ptr(), zd.data_ptr(), zd.data_ptr(),
                                      abd.data_ptr(), _lib.stream_ptr()))   # in place
    assert torch.equal(zd.cpu(), want)


# ------------------------------------------------------------------------------------------ norms
def test_rmsnorm_and_layernorm_rows(dev, lib):
    from multimodal_diffusion_b200.avdiff.models.mmdt import RMSNorm
    x = randn(5, 2, 133, 512) * 7
    x[0, 3] = 0.0                                         # a null-branch prompt row: must give exactly 0, not NaN
    scale = 1.0 + 0.1 * randn(6, 512)
    m = RMSNorm(512).to(dev)
    with torch.no_grad():
        m.scale.copy_(scale)
        got = m(x.to(dev)).cpu()
    want = orc.rmsnorm(x, scale)
    assert rel_l2(got, want) < 1e-6 and float(got[0, 3].abs().max()) == 0.0
    from multimodal_diffusion_b200 import _lib
    y, w, b = randn(7, 300, 512) * 3, 1 + 0.1 * randn(8, 512), 0.1 * randn(9, 512)
    out = torch.empty(300, 512, device=dev)
    yd, wd, bd = y.to(dev), w.to(dev), b.to(dev)
    _lib.check(lib.avd_layernorm_act(yd.data_ptr(), wd.data_ptr(), bd.data_ptr(), 300, 512, 1e-5, 1, out.data_ptr(),
                                     None, 0, _lib.stream_ptr()))
    assert rel_l2(out.cpu(), orc.gelu_erf(orc.layernorm(y, w, b))) < 2e-6


# ------------------------------------------------------------------------------------------ GEMM
def _gemm_case(dev, M, N, K, precision, **kw):
    from multimodal_diffusion_b200 import _lib, packing
    prec = _lib.precision_code(precision)
    A, W, bias = randn(1, M, K), randn(2, N, K) / math.sqrt(K), randn(3, N)
    Ad, Wd = A.to(dev), W.to(dev)
    xa, ssq = packing.rows_prepare(Ad, prec)
    Wp = packing.pack_linear_weight(Wd, prec)
    out = torch.zeros(M, N, device=dev)
    packing.gemm(xa, Wp, M, N, K, prec, bias=bias.to(dev), out_f32=out, **kw)
    torch.cuda.synchronize()
    if precision == "bf16":
        Ar, Wr = A.bfloat16().double(), W.bfloat16().double()
    else:
        Ar, Wr = A.double(), W.double()
    return out.cpu(), Ar @ Wr.t() + bias.double(), ssq.cpu(), A


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (300, 512, 512), (2128, 1536, 512), (266, 512, 2048),
                                   (74, 32, 512), (133, 256, 64)])
def test_gemm_bf16_exact_products(dev, M, N, K):
    got, want, ssq, A = _gemm_case(dev, M, N, K, "bf16")
    assert rel_l2(got, want) < 2e-6          # bf16 x bf16 products are exact in fp32; only accumulation order differs
    assert rel_l2(ssq.sum(-1), (A.double() ** 2).sum(-1)) < 1e-6


@pytest.mark.parametrize("precision,tol", [("bf16x3", 3e-5), ("fp32", 2e-6)])
@pytest.mark.parametrize("M,N,K", [(300, 512, 512), (266, 512, 2048)])
def test_gemm_split_precision(dev, M, N, K, precision, tol):
    got, want, _, _ = _gemm_case(dev, M, N, K, precision)
    assert rel_l2(got, want) < tol


@pytest.mark.parametrize("M,N,K", [(2128, 1536, 512), (300, 512, 512), (266, 512, 2048), (2128, 2048, 512)])
def test_gemm_cta_pair(dev, M, N, K):
    """The cta_group::2 variant (two SMs on one 256-row tile, opt-in through the gemm_pair option) against the same
    exact bf16 products, and bit-identical to the single-CTA kernel (same k order per accumulator)."""
    from multimodal_diffusion_b200 import _lib
    with _lib.option("gemm_pair", 2):
        got, want, _, _ = _gemm_case(dev, M, N, K, "bf16")
    with _lib.option("gemm_pair", 0):
        base, _, _, _ = _gemm_case(dev, M, N, K, "bf16")
    assert rel_l2(got, want) < 2e-6
    assert torch.equal(got, base)


def test_gemm_fused_epilogue(dev):
    """RMSNorm-as-row-scale prologue, bias, residual, GELU(erf), bf16-segment output and row ssq of the result."""
    from multimodal_diffusion_b200 import _lib, packing
    M, N, K = 266, 512, 512
    prec = _lib.AVD_PREC_FP32
    x, W, bias, resid = randn(1, M, K) * 4, randn(2, N, K) / math.sqrt(K), randn(3, N), randn(4, M, N)
    x[5] = 0.0
    scale = 1 + 0.1 * randn(5, K)
    xd = x.to(dev)
    xa, ssq = packing.rows_prepare(xd, prec)
    Wp = packing.pack_linear_weight(W.to(dev), prec, scale.to(dev))
    out = torch.empty(M, N, device=dev)
    out_b = torch.empty(3, M, N, device=dev, dtype=torch.bfloat16)
    ssq_o = torch.empty(M, N // 64, device=dev)
    packing.gemm(xa, Wp, M, N, K, prec, bias=bias.to(dev), ssq_in=ssq, rms_dim=K, rms_eps=1e-6, resid=resid.to(dev),
                 act=1, out_f32=out, out_bf16=out_b, out_precision=prec, ssq_out=ssq_o)
    want = orc.gelu_erf((orc.rmsnorm(x, scale).double() @ W.double().t() + bias.double() + resid.double()).float())
    assert rel_l2(out.cpu(), want) < 3e-6
    assert rel_l2(out_b.float().sum(0).cpu(), out.cpu()) < 1e-6           # hi + mid + lo reassembles the fp32 value
    assert rel_l2(ssq_o.sum(-1).cpu(), (out.cpu().double() ** 2).sum(-1)) < 1e-5


# ------------------------------------------------------------------------------------------ attention
def _attn_ref(qkv, n_seq, N, H, hd):
    d = H * hd
    q, k, v = qkv.double().view(n_seq, N, 3, H, hd).permute(2, 0, 3, 1, 4)
    p = torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(hd), dim=-1)
    return (p @ v).permute(0, 2, 1, 3).reshape(n_seq * N, d)


@pytest.mark.parametrize("n_seq,N", [(2, 133), (1, 128), (3, 445), (2, 16), (1, 300), (2, 129), (2, 136), (1, 137),
                                      (1, 192)])
def test_attention_tcgen05(dev, lib, n_seq, N):
    from multimodal_diffusion_b200 import _lib
    H, hd = 8, 64
    qkv = (randn(1, n_seq * N, 3 * H * hd) * 1.5).bfloat16()
    out = torch.zeros(1, n_seq * N, H * hd, device=dev, dtype=torch.bfloat16)
    qd = qkv.to(dev)
    _lib.check(lib.avd_attention(qd.data_ptr(), 0, out.data_ptr(), _lib.AVD_PREC_BF16, n_seq, N, H, hd,
                                 _lib.stream_ptr()))
    torch.cuda.synchronize()
    assert rel_l2(out[0].float().cpu(), _attn_ref(qkv.float(), n_seq, N, H, hd)) < 8e-3   # P and O rounded to bf16


@pytest.mark.parametrize("n_seq,N", [(2, 133), (1, 445)])
def test_attention_fp32(dev, lib, n_seq, N):
    from multimodal_diffusion_b200 import _lib
    H, hd = 8, 64
    qkv = randn(1, n_seq * N, 3 * H * hd) * 1.5
    out = torch.zeros(3, n_seq * N, H * hd, device=dev, dtype=torch.bfloat16)
    qd = qkv.to(dev)
    _lib.check(lib.avd_attention(qd.data_ptr(), 1, out.data_ptr(), _lib.AVD_PREC_FP32, n_seq, N, H, hd,
                                 _lib.stream_ptr()))
    assert rel_l2(out.float().sum(0).cpu(), _attn_ref(qkv, n_seq, N, H, hd)) < 2e-6


def _attn_ref_masked(qkv, n_seq, N, H, hd, mask):
    d = H * hd
    q, k, v = qkv.double().view(n_seq, N, 3, H, hd).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2) / math.sqrt(hd)).masked_fill(mask[:, None, None, :], float("-inf"))
    return (torch.softmax(s, dim=-1) @ v).permute(0, 2, 1, 3).reshape(n_seq * N, d)


@pytest.mark.parametrize("n_seq,N", [(3, 133), (2, 128), (2, 445), (2, 40)])
def test_attention_key_padding_mask(dev, lib, n_seq, N):
    """nn.MultiheadAttention's key_padding_mask (mmdt.py:58-60; §8f N4) on both attention kernels: trailing padding of
    different lengths per sequence plus scattered padded keys (a whole 32-key chunk among them)."""
    from multimodal_diffusion_b200 import _lib, packing
    H, hd = 8, 64
    gen = torch.Generator().manual_seed(11)
    mask = torch.rand(n_seq, N, generator=gen) < 0.15
    for b in range(n_seq):
        mask[b, N - 1 - 7 * b:] = True                  # ragged tails
    if N > 100:
        mask[0, 32:64] = True                           # one chunk fully padded
    mask[:, 0] = False                                  # every row keeps at least one key
    bits = packing.pack_key_padding_mask(mask.to(dev))
    qkv = (randn(2, n_seq * N, 3 * H * hd) * 1.5)
    want = _attn_ref_masked(qkv.bfloat16().float(), n_seq, N, H, hd, mask)
    out = torch.zeros(1, n_seq * N, H * hd, device=dev, dtype=torch.bfloat16)
    qd = qkv.bfloat16().to(dev)
    _lib.check(lib.avd_attention_masked(qd.data_ptr(), 0, out.data_ptr(), _lib.AVD_PREC_BF16, n_seq, N, H, hd,
                                        bits.data_ptr(), _lib.stream_ptr()))
    assert rel_l2(out[0].float().cpu(), want) < 8e-3
    out32 = torch.zeros(3, n_seq * N, H * hd, device=dev, dtype=torch.bfloat16)
    q32 = qkv.to(dev)
    _lib.check(lib.avd_attention_masked(q32.data_ptr(), 1, out32.data_ptr(), _lib.AVD_PREC_FP32, n_seq, N, H, hd,
                                        bits.data_ptr(), _lib.stream_ptr()))
    assert rel_l2(out32.float().sum(0).cpu(), _attn_ref_masked(qkv, n_seq, N, H, hd, mask)) < 3e-6


# ------------------------------------------------------------------------------------------ streaming stitchers (§8f N2)
def test_crossfade_audio_bit_exact(dev):
    """avd_crossfade_audio through the stream_infer mirror: reference fixtures, then seeded sizes against the oracle
    (vector path hop % 4 == 0, scalar path, gaps, one window, resident CUDA tensors)."""
    from oracle import stream_oracle as so
    from multimodal_diffusion_b200.avdiff.models.infer import stream_infer as si
    g = golden("stream")
    win, hop = (int(v) for v in g["a_win_hop"])
    assert np.array_equal(si.crossfade_audio(g["a_gen"], 1000, hop, win, 0.25), g["a_fade"])
    assert np.array_equal(si.crossfade_audio(g["a_gen"], 1000, hop, win, 0.0), g["a_rect"])
    assert np.array_equal(si.crossfade_audio(g["a_gen2"], 1000, 333, 1001, 0.1), g["a_odd"])
    assert np.array_equal(si.crossfade_audio(g["a_gen2"], 1000, 1200, 1001, 0.05), g["a_gap"])
    assert np.array_equal(si.crossfade_audio(g["a_gen2"], 1000, 500, 1001, 0.7), g["a_overlapfade"])
    rng = np.random.default_rng(5)
    for n, L, h, fade_s in ((1, 4096, 1024, 0.01), (9, 48000, 16000, 0.25), (17, 4000, 500, 0.02), (4, 999, 7, 0.0),
                            (3, 64, 64, 0.0), (6, 2048, 4096, 0.03)):
        ch = (rng.standard_normal((n, L)) * 3).astype(np.float32)
        want = so.crossfade_audio(ch, 16000, h, L, fade_s)
        assert np.array_equal(si.crossfade_audio(ch, 16000, h, L, fade_s), want), (n, L, h)
        got = si.crossfade_audio(torch.from_numpy(ch).to(dev), 16000, h, L, fade_s)
        assert got.is_cuda and np.array_equal(got.cpu().numpy(), want)
    with pytest.raises(ValueError):
        si.crossfade_audio(np.zeros((2, 100), np.float32), 16000, 50, 100, 1.0)      # fade longer than the window


def test_overlap_add_hann_bit_exact(dev):
    """ops.overlap_add_1d(apply_hann=True) (§8f N4) through avd_overlap_add_win against the reference's outputs."""
    from multimodal_diffusion_b200.avdiff.utils import ops
    g = golden("stream")
    w = torch.from_numpy(g["hann_in"]).to(dev)
    for s in (3, 5, 7):
        assert np.array_equal(ops.overlap_add_1d(w, stride=s, apply_hann=True).cpu().numpy(), g[f"hann_s{s}"])
    w2 = torch.from_numpy(g["hann_in2"]).to(dev)
    assert np.array_equal(ops.overlap_add_1d(w2, stride=4, apply_hann=True).cpu().numpy(), g["hann2_s4"])
    x = randn(31, 3, 20, 12, 8)                                   # windows on dim 2 (not the default -2)
    want = orc.overlap_add_1d(x, stride=5, dim_windows=2, apply_hann=True)
    assert torch.equal(ops.overlap_add_1d(x.to(dev), stride=5, dim_windows=2, apply_hann=True).cpu(), want)


def test_crossfade_video_bit_exact(dev):
    from oracle import stream_oracle as so
    from multimodal_diffusion_b200.avdiff.models.infer import stream_infer as si
    g = golden("stream")
    vwin, vhop = (int(v) for v in g["v_win_hop"])
    assert np.array_equal(si.crossfade_video(g["v_gen"], vhop, vwin, 4), g["v_fade"])
    assert np.array_equal(si.crossfade_video(g["v_gen"], vhop, vwin, 0), g["v_nofade"])
    assert np.array_equal(si.crossfade_video(g["v_gen2"], 5, 7, 3), g["v_odd"])
    rng = np.random.default_rng(6)
    for n, T, H, W, h, fade in ((1, 48, 16, 16, 16, 4), (7, 48, 128, 128, 16, 4), (5, 12, 9, 7, 5, 0), (3, 8, 4, 4, 20, 2)):
        ch = rng.integers(0, 256, size=(n, T, H, W, 3), dtype=np.uint8)
        want = so.crossfade_video(ch, h, T, fade)
        assert np.array_equal(si.crossfade_video(ch, h, T, fade), want), (n, T, H, W, h)
        got = si.crossfade_video(torch.from_numpy(ch).to(dev), h, T, fade)
        assert got.is_cuda and np.array_equal(got.cpu().numpy(), want)


def test_stream_kernels_stay_in_bounds(dev, lib):
    """Guard bands around the outputs of the stitching kernels (odd sizes: scalar and vector paths, grid tails): no
    byte outside [(N-1)*hop + L] may change.  (compute-sanitizer is not available on the pool.)"""
    from multimodal_diffusion_b200 import _lib
    G = 1024
    for n, L, hop in ((3, 1001, 333), (5, 4096, 1024), (2, 64, 200), (7, 48000, 16000)):
        T_out = (n - 1) * hop + L
        ch = torch.randn(n, L, device=dev)
        w = torch.rand(L, device=dev)
        buf = torch.full((T_out + 2 * G,), 7.0, device=dev)
        _lib.check(lib.avd_crossfade_audio(ch.data_ptr(), w.data_ptr(), buf[G:].data_ptr(), n, L, hop, _lib.stream_ptr()))
        assert bool((buf[:G] == 7.0).all()) and bool((buf[G + T_out:] == 7.0).all())
        rows = 3
        chr_ = torch.randn(rows, n, L, device=dev)
        buf = torch.full((rows * T_out + 2 * G,), 7.0, device=dev)
        _lib.check(lib.avd_overlap_add_win(chr_.data_ptr(), w.data_ptr(), buf[G:].data_ptr(), rows, n, L, hop, 1e-8,
                                           _lib.stream_ptr()))
        assert bool((buf[:G] == 7.0).all()) and bool((buf[G + rows * T_out:] == 7.0).all())
        assert not bool((buf[G:G + rows * T_out] == 7.0).all())
    for n, T, F, hop in ((3, 7, 45, 5), (4, 12, 48 * 3, 4), (2, 5, 16 * 16 * 3, 9)):
        T_out = (n - 1) * hop + T
        ch = torch.randint(0, 256, (n, T, F), device=dev, dtype=torch.uint8)
        w = torch.rand(T, device=dev)
        buf = torch.full((T_out * F + 2 * G,), 9, device=dev, dtype=torch.uint8)
        _lib.check(lib.avd_crossfade_video(ch.data_ptr(), w.data_ptr(), buf[G:].data_ptr(), n, T, F, hop, _lib.stream_ptr()))
        assert bool((buf[:G] == 9).all()) and bool((buf[G + T_out * F:] == 9).all())
