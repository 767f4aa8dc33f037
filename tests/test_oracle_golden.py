"""CPU: the oracle (oracle/avdiff_oracle.py) against fixtures produced by the unmodified reference
(tests/golden/make_golden.py), plus the one numerical pin the reference's own tests hold
(tests/test_shapes.py:26-36, tube patch/unpatch round trip)."""
import numpy as np
import torch

from _common import golden, orc, randn, rel_l2, seeded_modules, weights_dict


def T(a):
    return torch.from_numpy(np.asarray(a))


def test_reference_own_pin_tube_roundtrip():
    torch.manual_seed(0)                                  # tests/test_shapes.py:24
    z = torch.randn(2, 8, 12, 16, 16)
    tok = orc.tube_patch_video(z, 2, 4, 4)
    assert tok.shape == (2, 96, 256)
    assert torch.allclose(z, orc.tube_unpatch_video(tok, 8, 12, 16, 16, 2, 4, 4), atol=1e-6)


def test_tokenizers_bit_exact():
    g = golden("tokenizers")
    z = randn(int(g["tube_z_seed"]), 2, 8, 12, 16, 16)
    tok = orc.tube_patch_video(z, 2, 4, 4)
    assert np.array_equal(tok[:, ::7, ::5].numpy(), g["tube_tok_sha"])
    assert bool(g["tube_roundtrip_equal"])
    zs = T(g["tube_small_z"])
    for key, (t, h, w) in {"tube_small_tok_t2h3w2": (2, 3, 2), "tube_small_tok_t1h2w1": (1, 2, 1)}.items():
        tk = orc.tube_patch_video(zs, t, h, w)
        assert np.array_equal(tk.numpy(), g[key])
        assert torch.equal(orc.tube_unpatch_video(tk, 3, 4, 6, 8, t, h, w), zs)
        idx = orc.tube_patch_index_numpy(1, 3, 4, 6, 8, t, h, w)
        assert np.array_equal(zs.numpy().reshape(-1)[idx], g[key].reshape(-1))
    za = T(g["aud_z"])
    assert np.array_equal(orc.chunk_1d(za, 4, 4).numpy(), g["aud_windows_l4s4"])
    assert np.array_equal(orc.latents_to_tokens_audio(za, 4, 4).numpy(), g["aud_tok_l4s4"])
    assert np.array_equal(g["aud_tok_l4s4"], g["aud_tok_l4s4_cls"])
    assert np.array_equal(orc.chunk_1d(za, 6, 2).numpy(), g["aud_windows_l6s2"])
    assert np.array_equal(orc.latents_to_tokens_audio(za, 6, 2).numpy(), g["aud_tok_l6s2"])
    tk = T(g["aud_untok_in"])
    u150 = orc.tokens_to_latents_audio(tk, 8, 4, 150, 4)
    assert np.array_equal(u150.numpy(), g["aud_untok_l4s4_F150"])
    assert float(u150[..., 148:].abs().max()) == 0.0       # the tail frames get eps = 0
    assert np.array_equal(orc.tokens_to_latents_audio(tk, 8, 4, 140, 4).numpy(), g["aud_untok_l4s4_F140"])
    assert np.array_equal(orc.tokens_to_latents_audio(T(g["aud_untok2_in"]), 8, 6, 150, 2).numpy(),
                          g["aud_untok_l6s2_F150"])
    w = T(g["ola_in"])
    for s in (3, 5, 7):
        assert np.array_equal(orc.overlap_add_1d(w, stride=s).numpy(), g[f"ola_s{s}"])
    assert np.array_equal(orc.chunk_1d(randn(6, 2, 3), 4, 4).numpy(), g["chunk_short"])
    assert orc.num_chunks(150, 4, 4) == 37 and orc.num_chunks(3, 4, 4) == 1


def test_schedules_bit_exact():
    g = golden("schedules")
    for kind in ("cosine", "linear", "sigmoid"):
        b = orc.make_beta_schedule(1000, kind, 1e-4, 0.02)
        assert np.array_equal(b.numpy(), g[f"betas_{kind}"])
        assert np.array_equal(orc.alphas_cumprod_from_betas(b)[1].numpy(), g[f"alpha_bar_{kind}"])
    for S in (10, 25, 50, 60):
        assert np.array_equal(orc.make_sampling_schedule(1000, S).numpy(), g[f"sched_{S}"])
    ts = orc.make_sampling_schedule(1000, 50)
    assert ts[0] == 999 and ts[-1] == -1 and int(ts[1]) == 979
    assert np.array_equal(orc.timestep_embedding(ts[:-1], 256).numpy(), g["temb_steps_256"])
    assert np.array_equal(orc.timestep_embedding(torch.zeros(1, dtype=torch.long), 256).numpy(), g["temb_zero_256"])
    assert np.array_equal(orc.timestep_embedding(torch.tensor([0, 3, 999]), 7).numpy(), g["temb_odd_7"])
    ab = T(g["alpha_bar_cosine"])
    x, e = T(g["ddim_x"]), T(g["ddim_eps"])
    for name, (tn, tp) in {"first": (999, 979), "mid": (499, 479), "last": (19, -1)}.items():
        out = orc.ddim_step(x, torch.full((3,), tn), torch.full((3,), tp), e, ab)
        assert np.array_equal(out.numpy(), g[f"ddim_{name}"])
    out = orc.ddim_step(x, torch.tensor([999, 499, 19]), torch.tensor([979, 479, -1]), e, ab)
    assert np.array_equal(out.numpy(), g["ddim_mixed_t"])


def test_seeded_init_matches_reference_checksums():
    g = golden("mvp_forward")
    for name, mod in zip(("adapt_v", "adapt_a", "core", "head"), seeded_modules()):
        sd = mod.state_dict()
        keys = [str(k) for k in g[f"ck_{name}_keys"]]
        assert list(sd.keys()) == keys, f"{name}: state_dict keys differ from the reference's"
        vals = np.array([[float(v.double().sum()), float(v.double().abs().sum())] for v in sd.values()])
        assert np.array_equal(vals, g[f"ck_{name}_vals"]), f"{name}: seeded init differs from the reference's"


def test_core_and_head_match_reference():
    g = golden("mvp_forward")
    w = weights_dict()
    X = T(g["X"])
    with torch.no_grad():
        b0 = orc.block_forward(w["core"], 0, X, 8)
        assert rel_l2(b0, g["block0_out"]) < 2e-6
        h_c = orc.mmdit_forward(w["core"], X, 8, 8)
        assert rel_l2(h_c, g["h_cond"]) < 5e-6
        X_null = torch.cat([torch.zeros_like(X[:, :96]), X[:, 96:]], dim=1)
        h_n = orc.mmdit_forward(w["core"], X_null, 8, 8)
        assert rel_l2(h_n, g["h_null"]) < 5e-6
        assert rel_l2(orc.head_forward(w["head"], T(g["h_cond"])[:, 96:], "audio"), g["eps_cond_audio"]) < 2e-6
        assert rel_l2(orc.head_forward(w["head"], T(g["h_null"])[:, 96:], "audio"), g["eps_null_audio"]) < 2e-6
        assert rel_l2(orc.head_forward(w["head"], T(g["h_cond"])[:, :96], "video"), g["eps_cond_video"]) < 2e-6


def test_joint_sequence_matches_reference():
    g = golden("mvp_forward")
    w = weights_dict()
    cfg = orc.HotPathConfig()
    z_v, z_a = randn(11, 2, 8, 12, 16, 16), randn(12, 2, 8, 150)
    X, Nv = orc.build_joint_sequence(cfg, w, z_v, z_a, torch.zeros(2, dtype=torch.long), torch.full((2,), 999))
    assert Nv == 96 and rel_l2(X, g["X"]) < 1e-6


def _check_sampling(name, target):
    g = golden(f"sample_{name}")
    w = weights_dict()
    cfg = orc.HotPathConfig(sampler_steps=50, guidance=3.5)
    assert np.array_equal(cfg.timesteps()[:-1].numpy(), g["t_now"])
    assert np.array_equal(cfg.timesteps()[1:].numpy(), g["t_prev"])
    rec = []
    z = orc.sample_loop(cfg, w, target, T(g["z_prompt"]), T(g["z_init"]), record=rec, record_steps=(0, 1, 49))
    by = {r["step"]: r for r in rec}
    for s in (0, 1, 49):
        assert rel_l2(by[s]["eps_hat_lat"], g[f"eps_hat_lat_{s}"]) < 1e-4, s
        assert rel_l2(by[s]["z_prev"], g[f"z_prev_{s}"]) < 1e-4, s
    assert rel_l2(z, g["z_final"]) < 1e-3
    return z, g


def test_sampling_v2a_matches_reference():
    z, g = _check_sampling("v2a", "audio")
    # audio frames 148,149 never see the model; DDIM with eps = 0 only rescales them
    assert z.shape == (1, 8, 150)


def test_sampling_a2v_matches_reference():
    _check_sampling("a2v", "video")


def test_batched_oracle_equals_per_clip_runs():
    """The reference hard-codes B = 1 (sample_clip.py:281); the oracle's batched loop must reproduce per-clip runs."""
    w = weights_dict()
    cfg = orc.HotPathConfig(sampler_steps=3, guidance=3.5)
    zp, zi = randn(31, 2, 8, 12, 16, 16), randn(32, 2, 8, 150)
    zb = orc.sample_loop(cfg, w, "audio", zp, zi)
    for b in range(2):
        z1 = orc.sample_loop(cfg, w, "audio", zp[b:b + 1], zi[b:b + 1])
        assert rel_l2(zb[b:b + 1], z1) < 1e-5


# ------------------------------------------------------------------------------------------ streaming driver (§8f N2)
def test_stream_oracle_and_host_split_match_reference():
    """oracle/stream_oracle.py and the product's host-side window splitting against outputs of the reference's own
    stream_infer functions (tests/golden/stream.npz, generated by make_golden_stream.py): bit-exact."""
    from oracle import stream_oracle as so
    from multimodal_diffusion_b200.avdiff.models.infer import stream_infer as si
    g = golden("stream")
    win, hop = (int(v) for v in g["a_win_hop"])
    for split_a, split_v in ((so.split_audio, so.split_frames), (si.split_audio_into_windows, si.split_frames_into_windows)):
        ch, w_, h_ = split_a(g["a_y"], 1000, 3.0, 1.0)
        assert (w_, h_) == (win, hop) and np.array_equal(ch, g["a_chunks"])
        assert np.array_equal(split_a(g["a_y"][:2500], 1000, 3.0, 1.0)[0], g["a_short"])
        vch, vw, vh = split_v(g["v_frames"], 16, 3.0, 1.0)
        assert (vw, vh) == tuple(int(v) for v in g["v_win_hop"]) and np.array_equal(vch, g["v_chunks"])
        assert np.array_equal(split_v(g["v_frames"][:30], 16, 3.0, 1.0)[0], g["v_short"])
    assert np.array_equal(so.crossfade_audio(g["a_gen"], 1000, hop, win, 0.25), g["a_fade"])
    assert np.array_equal(so.crossfade_audio(g["a_gen"], 1000, hop, win, 0.0), g["a_rect"])
    assert np.array_equal(so.crossfade_audio(g["a_gen2"], 1000, 333, 1001, 0.1), g["a_odd"])
    assert np.array_equal(so.crossfade_audio(g["a_gen2"], 1000, 1200, 1001, 0.05), g["a_gap"])
    assert np.array_equal(so.crossfade_audio(g["a_gen2"], 1000, 500, 1001, 0.7), g["a_overlapfade"])
    vwin, vhop = (int(v) for v in g["v_win_hop"])
    assert np.array_equal(so.crossfade_video(g["v_gen"], vhop, vwin, 4), g["v_fade"])
    assert np.array_equal(so.crossfade_video(g["v_gen"], vhop, vwin, 0), g["v_nofade"])
    assert np.array_equal(so.crossfade_video(g["v_gen2"], 5, 7, 3), g["v_odd"])
    assert np.array_equal(si.audio_fade_window(3000, 250), so.audio_window(3000, 250))
    assert np.array_equal(si.video_fade_window(48, 4), so.video_window(48, 4))
    for s in (3, 5, 7):
        assert np.array_equal(orc.overlap_add_1d(torch.from_numpy(g["hann_in"]), stride=s, apply_hann=True).numpy(),
                              g[f"hann_s{s}"])
    assert np.array_equal(orc.overlap_add_1d(torch.from_numpy(g["hann_in2"]), stride=4, apply_hann=True).numpy(),
                          g["hann2_s4"])
    # hop larger than the window and a hop that overshoots the end (the while-loop's exit conditions)
    for total, w_, h_ in ((10, 3, 8), (10, 3, 20), (9, 3, 3), (10, 3, 3)):
        assert si._window_starts(total, w_, h_) == so.window_starts(total, w_, h_)
