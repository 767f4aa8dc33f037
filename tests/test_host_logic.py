"""CPU: host-side logic of the product (config parsing, derived shapes, FLOP model, clip sharding incl. a
world_size-2 gloo run, error behaviour of the mirror ops before any launch)."""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

from _common import golden

ROOT = Path(__file__).resolve().parent.parent


def test_hot_path_config_shapes_and_flops():
    from multimodal_diffusion_b200.sampler import HotPathConfig
    c = HotPathConfig()
    assert (c.Nv, c.Na, c.Dv, c.Da) == (96, 37, 256, 32)
    # per clip (50 steps) V2A 7.044e11, A2V 7.166e11 FLOP
    assert abs(c.flops_per_clip("audio") / 7.044e11 - 1) < 2e-3
    assert abs(c.flops_per_clip("video") / 7.166e11 - 1) < 2e-3
    long = HotPathConfig(Fa=500, video_latent=(40, 16, 16))
    assert long.Nv + long.Na == 445 and abs(long.flops_per_clip("video") / 2.625e12 - 1) < 2e-3


def test_config_from_yaml_dict():
    from multimodal_diffusion_b200.sampler import HotPathConfig
    cfg = {
        "tokenizer": {"width": 512, "video": {"tube": {"t": 2, "h": 4, "w": 4}}, "audio": {"chunk": {"length": 4, "stride": 4}}},
        "embeddings": {"timestep_dim": 256},
        "model": {"core": {"d_model": 512, "n_layers": 8, "n_heads": 8, "mlp_ratio": 4.0},
                  "heads": {"video": {"out_dim": 256, "hidden_dim": 512}, "audio": {"out_dim": 32, "hidden_dim": 512}}},
        "video": {"fps": 16, "size": [128, 128], "latent": {"channels": 8, "t_down": 4, "s_down": 8}},
        "audio": {"sr": 16000, "latent": {"channels": 8, "frames_per_clip": 150}},
        "data": {"clip_seconds": 3.0},
        "diffusion": {"video": {"steps": 1000, "sampler_steps": 60, "schedule": "cosine", "min_beta": 1e-4, "max_beta": 0.02},
                      "audio": {"steps": 1000, "sampler_steps": 50, "schedule": "cosine", "min_beta": 1e-4, "max_beta": 0.02}},
        "sampling": {"ddim_eta": 0.0, "guidance_scale": {"video": 3.5, "audio": 0.0}},
    }
    hp = HotPathConfig.from_yaml_dict(cfg, "video")
    assert hp.video_latent == (12, 16, 16) and hp.sampler_steps == 60 and hp.guidance == 3.5 and hp.tube == (2, 4, 4)
    assert HotPathConfig.from_yaml_dict(cfg, "video", sampler_steps=50).sampler_steps == 50


def test_host_tables_bit_exact_vs_reference():
    from multimodal_diffusion_b200.avdiff.utils import schedule_utils as su
    from multimodal_diffusion_b200.avdiff.models.schedules import ModalitySchedule, build_schedules_from_config
    g = golden("schedules")
    for kind in ("cosine", "linear", "sigmoid"):
        b = su.make_beta_schedule(1000, kind, 1e-4, 0.02)
        assert np.array_equal(b.numpy(), g[f"betas_{kind}"])
        assert np.array_equal(su.alphas_cumprod_from_betas(b)[1].numpy(), g[f"alpha_bar_{kind}"])
    for S in (10, 25, 50, 60):
        assert np.array_equal(su.make_sampling_schedule(1000, S).numpy(), g[f"sched_{S}"])
    assert np.array_equal(su.timestep_embedding_host(su.make_sampling_schedule(1000, 50)[:-1], 256).numpy(),
                          g["temb_steps_256"])
    with pytest.raises(ValueError):
        su.make_beta_schedule(10, "bogus")
    sch = ModalitySchedule.make(kind="cosine", steps=1000)
    assert np.array_equal(sch.alphas_cumprod.numpy(), g["alpha_bar_cosine"])
    assert np.array_equal(sch.make_sampling_schedule(25).numpy(), g["sched_25"])
    both = build_schedules_from_config({"diffusion": {"video": {}, "audio": {"schedule": "linear"}}})
    assert both["audio"].kind == "linear" and np.array_equal(both["audio"].betas.numpy(), g["betas_linear"])


def test_shard_range_partitions_exactly():
    from multimodal_diffusion_b200.sampler import shard_range
    for n in (0, 1, 7, 8, 64, 65):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert [shard_range(64, r, 8) for r in (0, 7)] == [(0, 8), (56, 64)]


def test_clip_noise_is_shard_invariant():
    from multimodal_diffusion_b200.sampler import clip_noise
    a = torch.stack([clip_noise((8, 150), i, 1000) for i in range(16)])
    b = torch.cat([torch.stack([clip_noise((8, 150), i, 1000) for i in range(lo, hi)]) for lo, hi in ((0, 8), (8, 16))])
    assert torch.equal(a, b) and not torch.equal(a[0], a[1])


def test_mirror_errors_before_any_launch():
    """Argument errors are raised on the host exactly where the reference raises them (no GPU needed)."""
    from multimodal_diffusion_b200 import _lib
    from multimodal_diffusion_b200.avdiff.utils import ops
    from multimodal_diffusion_b200.avdiff.models.tokenizers import AudioTokenizer, VideoTokenizer
    with pytest.raises(AssertionError, match="tube sizes"):
        ops.tube_patch_video(torch.zeros(1, 2, 3, 4, 4), 2, 2, 2)                  # ops.py:113
    with pytest.raises(AssertionError, match="token width"):
        ops.tube_unpatch_video(torch.zeros(1, 4, 9), 2, 2, 4, 4, 1, 2, 2)           # ops.py:137
    with pytest.raises(AssertionError, match="token count"):
        ops.tube_unpatch_video(torch.zeros(1, 3, 8), 2, 2, 4, 4, 1, 2, 2)           # ops.py:139
    with pytest.raises(AssertionError):
        AudioTokenizer().decode(torch.zeros(1, 3, 7), 8, 150)                        # tokenizers.py:96
    assert VideoTokenizer().token_dim(8) == 256 and AudioTokenizer().token_dim(8) == 32
    with pytest.raises(_lib.AvdError, match="CUDA"):                                 # no CPU path: loud failure
        ops.tube_patch_video(torch.zeros(1, 2, 2, 4, 4), 2, 2, 2)
    with pytest.raises(ValueError):
        _lib.precision_code("fp8")


GLOO_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from multimodal_diffusion_b200.sampler import shard_range, clip_noise
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
n = 6
lo, hi = shard_range(n, rank, world)
# stand-in for the per-rank sampler output: a deterministic function of the clip's own seeded noise only
mine = torch.stack([clip_noise((8, 150), i, 2000) * 2.0 + 1.0 for i in range(lo, hi)])
gathered = [torch.empty_like(mine) for _ in range(world)]
dist.all_gather(gathered, mine)                      # the path's only collective: gather final latents
full = torch.cat(gathered)
want = torch.stack([clip_noise((8, 150), i, 2000) * 2.0 + 1.0 for i in range(n)])
assert torch.equal(full, want), "sharded + gathered result differs from the single-rank result"
t = torch.tensor([float(rank + 1)], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)             # bench.py's max-over-ranks timing reduction
assert t.item() == world
dist.barrier(); dist.destroy_process_group()
print("ok", rank)
'''


def test_two_rank_gloo_sharding(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(GLOO_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29731", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script), str(ROOT)], env=dict(env, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert all("ok" in o for o in outs)
