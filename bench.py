#!/usr/bin/env python3
"""bench.py -- clips/s of 50-step DDIM + CFG sampling at mvp.yaml shapes (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--direction v2a|a2v] [--batch B]

A "step" is ONE sampling call: 50 DDIM steps x (cond + null) MMDT forwards + head + fused CFG/DDIM update for a
batch of B clips -- the workload of BASELINE.json configs[1] (V2A, batch 8, 1 x B200) by default.  With N > 1
(torchrun, one rank per GPU) every rank samples its own B clips (clip-sharded, weak scaling, no data-path
collective) and the final latents are all-gathered over NCCL inside the timed step.

Prints ONE JSON line (rank 0).  `value` = whole-job clips/s with inputs resident in HBM; `e2e` = same through
the public API with pinned-host inputs (H2D + D2H inside the timed region); `roofline` = the tcgen05 GEMM kernel
family against the measured bf16 peak; `cpu_baseline` = the CPU oracle port timed on this host's cores.
`--impl reference` times the reference's CPU algorithm (the oracle port: the reference is pure PyTorch-CPU code,
DESIGN.md §6) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

SAMPLER_STEPS = 50
GUIDANCE = 3.5
_CPU_WEIGHTS = None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--direction", default="v2a", choices=["v2a", "a2v"])
    ap.add_argument("--batch", type=int, default=8, help="clips per GPU per sampling call")
    ap.add_argument("--core-precision", default="f16",
                    help="f16 (default: fp16 operands, fp32 accumulation; meets every tolerance of the hot-path spec), bf16, bf16x3, fp32")
    ap.add_argument("--head-precision", default="f16")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-steps", type=int, default=10, help="DDIM steps of the CPU-baseline sample")
    ap.add_argument("--chains", type=int, default=0, help="sequence groups on forked streams (library option; 0 = keep)")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the extra legs (strong-scaling A2V batch 64, long context, PyTorch-CUDA baseline)")
    ap.add_argument("--ref-budget-s", type=float, default=240.0,
                    help="--impl reference: wall-clock budget that decides between full 50-step calls and a bounded sample")
    return ap.parse_args()


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"bf16_tflops": float(d["bf16_tflops"]), "bf16_tflops_sustained": float(d["bf16_tflops_sustained"]),
                "hbm_gbs": float(d["hbm_gbs"]), "source": "measured"}
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md's clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------
# CPU side: the oracle port of the reference's algorithm (test infrastructure; here only as the timed baseline)
# ---------------------------------------------------------------------------------------------------
def cpu_sample_seconds(direction: str, batch: int, ddim_steps: int, threads: int) -> float:
    """Wall time of `ddim_steps` DDIM+CFG steps of the oracle for `batch` clips on the host cores."""
    import torch
    from _common import orc, randn, weights_dict
    torch.set_num_threads(threads)
    target = "audio" if direction == "v2a" else "video"
    global _CPU_WEIGHTS
    if _CPU_WEIGHTS is None:
        _CPU_WEIGHTS = weights_dict()
    w = _CPU_WEIGHTS
    cfg = orc.HotPathConfig(sampler_steps=SAMPLER_STEPS, guidance=GUIDANCE)
    zp = randn(1, batch, 8, 12, 16, 16) if target == "audio" else randn(1, batch, 8, 150)
    zi = randn(2, batch, 8, 150) if target == "audio" else randn(2, batch, 8, 12, 16, 16)
    ab, ts = cfg.alpha_bar(), cfg.timesteps()
    z = zi.clone()
    t0 = time.perf_counter()
    with torch.no_grad():
        for i in range(ddim_steps):
            t_now, t_prev = ts[i].repeat(batch), ts[i + 1].repeat(batch)
            z_v, z_a = (zp, z) if target == "audio" else (z, zp)
            _, _, e_hat = orc.eps_for_step(cfg, w, target, z_v, z_a, t_now)
            z = orc.ddim_step(z, t_now, t_prev, orc.untokenise(cfg, target, e_hat, z), ab)
    return time.perf_counter() - t0


def torch_cuda_seconds(direction: str, batch: int, ddim_steps: int, autocast_bf16: bool) -> float:
    """The same oracle loop with every tensor on cuda:current -- eager PyTorch (cuBLAS / library attention kernels) on the
    SAME GPU: SURVEY §8(d)'s "reference on same GPU" line.  fp32 eager (TF32 off, as the reference runs) or bf16 autocast."""
    import torch
    from _common import orc, randn, weights_dict
    dev = torch.device("cuda", torch.cuda.current_device())
    target = "audio" if direction == "v2a" else "video"
    w = {k: {n: t.to(dev) for n, t in sd.items()} for k, sd in weights_dict().items()}
    cfg = orc.HotPathConfig(sampler_steps=SAMPLER_STEPS, guidance=GUIDANCE)
    zp = (randn(1, batch, 8, 12, 16, 16) if target == "audio" else randn(1, batch, 8, 150)).to(dev)
    zi = (randn(2, batch, 8, 150) if target == "audio" else randn(2, batch, 8, 12, 16, 16)).to(dev)
    ab, ts = cfg.alpha_bar().to(dev), cfg.timesteps().to(dev)

    import math
    from oracle import avdiff_oracle as _orc_mod

    def mha_sdpa(x, w_in, b_in, w_out, b_out, n_heads, key_padding_mask=None):
        # nn.MultiheadAttention's fast path ends in the library attention kernel (mmdt.py:58-61); the explicit-math
        # softmax of the CPU port is replaced by F.scaled_dot_product_attention for this GPU baseline
        B, N, d = x.shape
        hd = d // n_heads
        q, k, v = (x @ w_in.t() + b_in).split(d, dim=-1)
        q, k, v = (t.reshape(B, N, n_heads, hd).transpose(1, 2) for t in (q, k, v))
        a = torch.nn.functional.scaled_dot_product_attention(q, k, v, scale=1.0 / math.sqrt(hd))
        return a.transpose(1, 2).reshape(B, N, d) @ w_out.t() + b_out

    def run(n):
        z = zi.clone()
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast_bf16):
            for i in range(n):
                t_now, t_prev = ts[i].repeat(batch), ts[i + 1].repeat(batch)
                z_v, z_a = (zp, z) if target == "audio" else (z, zp)
                _, _, e_hat = orc.eps_for_step(cfg, w, target, z_v, z_a, t_now)
                z = orc.ddim_step(z.float(), t_now, t_prev, orc.untokenise(cfg, target, e_hat.float(), z), ab)
        return z

    saved, tf32 = _orc_mod.mha, torch.backends.cuda.matmul.allow_tf32
    _orc_mod.mha = mha_sdpa
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        run(2)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(ddim_steps)
        e1.record()
        torch.cuda.synchronize()
    finally:
        _orc_mod.mha = saved
        torch.backends.cuda.matmul.allow_tf32 = tf32
    return e0.elapsed_time(e1) * 1e-3


def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def workload_name(direction: str, batch: int) -> str:
    shapes = "video prompt [B,8,12,16,16] -> audio latents [B,8,150]" if direction == "v2a" else \
        "audio prompt [B,8,150] -> video latents [B,8,12,16,16]"
    return (f"configs/{direction}.yaml {direction.upper()} sampling at mvp.yaml shapes ({shapes}), "
            f"{SAMPLER_STEPS}-step DDIM + CFG g={GUIDANCE}, batch {batch} per GPU, N=133 joint tokens")


def run_reference(args) -> None:
    """--impl reference: the reference's CPU algorithm (oracle port of sample_clip.py:317-389, fp32 torch-CPU, all host
    threads) on the same workload.  Each timed step is ONE full 50-step sampling call for the batch whenever
    (warmup + steps) such calls fit `--ref-budget-s` (no extrapolation); otherwise a bounded sample (a fixed number of
    the 50 DDIM steps, whose per-step cost is constant) and the value is scaled -- the line then says so
    (`sample_scale` > 1, and `ms_per_step` stays the MEASURED time of a step so it agrees with the wall clock)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    t_probe = cpu_sample_seconds(args.direction, args.batch, 1, threads)          # also the warm-up of torch's threads
    t_probe = min(t_probe, cpu_sample_seconds(args.direction, args.batch, 1, threads))
    calls = args.steps + max(args.warmup, 0)
    full = t_probe * SAMPLER_STEPS * calls <= args.ref_budget_s
    sub = SAMPLER_STEPS if full else max(1, min(SAMPLER_STEPS, int(args.ref_budget_s / max(calls * t_probe, 1e-9))))
    for _ in range(args.warmup):
        cpu_sample_seconds(args.direction, args.batch, sub, threads)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_sample_seconds(args.direction, args.batch, sub, threads)
    measured = t / args.steps                                   # seconds per timed step (as run)
    scale = SAMPLER_STEPS / sub
    per_call = measured * scale                                 # seconds for one full 50-step sampling call
    value = args.batch / per_call
    # the reference itself runs B = 1 sequentially (sample_clip.py:281): the same port at batch 1, a short sample
    t1 = cpu_sample_seconds(args.direction, 1, 5, threads)
    b1 = 1.0 / (t1 * SAMPLER_STEPS / 5)
    sample = (f"{'the full ' if full else ''}{sub} of {SAMPLER_STEPS} DDIM+CFG steps for the full batch of {args.batch} clips "
              f"per timed step{'' if full else f', scaled x{scale:g}'}; fp32 torch-CPU oracle port of "
              f"sample_clip.py:317-389 (the reference is pure PyTorch and does not travel to this box)")
    line = {"metric": "DDIM+CFG clips/sec (mvp.yaml)", "value": value, "unit": "clips/s", "impl": "reference",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": measured * 1e3,
            "ms_per_sampling_call": per_call * 1e3, "sample_scale": scale, "extrapolated": not full,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.direction, args.batch), "direction": args.direction,
                       "batch_per_gpu": args.batch, "sampler_steps": SAMPLER_STEPS},
            "cpu_baseline": {"value": value, "unit": "clips/s", "cores": threads, "kind": "port", "sample": sample,
                             "b1_sequential": {"value": b1, "unit": "clips/s",
                                               "sample": "5 of 50 DDIM+CFG steps at batch 1 (as the reference runs), scaled x10"}},
            "e2e": {"value": value, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------------------
def gemm_flops_per_call(cfg, target: str, batch: int) -> float:
    """Algorithmic FLOPs of all tcgen05-GEMM launches of one sampling call: per DDIM step, with
    M = 2B*N rows through the core and Mt = 2B*Nt rows through the head."""
    d, L, N = cfg.d_model, cfg.n_layers, cfg.Nv + cfg.Na
    dm = int(d * cfg.mlp_ratio)
    Nt, Dt = (cfg.Na, cfg.Da) if target == "audio" else (cfg.Nv, cfg.Dv)
    M, Mt, hid = 2 * batch * N, 2 * batch * Nt, cfg.head_hidden
    core = L * 2.0 * M * (3 * d * d + d * d + 2 * d * dm)
    head = 2.0 * Mt * (d * hid + 2 * hid * hid + hid * Dt)
    return cfg.sampler_steps * (core + head)


def attn_flops_per_call(cfg, batch: int) -> float:
    N = cfg.Nv + cfg.Na
    return cfg.sampler_steps * cfg.n_layers * 2 * batch * 4.0 * N * N * cfg.d_model


def ncu_traffic_bytes():
    """dram read + write bytes per GEMM launch from the committed `ncu --set full` capture of one block's four GEMM
    launches at the bench shapes (the newest profiles/r*_gemm_mvp_ncu_full_summary.csv, written by
    tools/round_artifacts.sh: cold-cache, serialised replays), or None."""
    import csv
    found = sorted((ROOT / "profiles").glob("r*_gemm_mvp_ncu_full_summary.csv"))
    if not found:
        return None
    f = found[-1]
    try:
        rows = list(csv.reader(f.open()))
        hdr, units = rows[0], rows[1]
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot = [float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]] for r in rows[2:] if len(r) > max(ir, iw)]
        return sum(tot) / len(tot) if tot else None
    except Exception:
        return None


def time_sampler_calls(fs, zp_d, zi_d, out_d, flush, iters: int, warm: int = 2) -> float:
    """Device-timed ms per sampling call of a captured FusedSampler (inputs resident, L2 flushed between calls)."""
    import torch
    for _ in range(warm):
        fs.sample(zp_d, zi_d, out=out_d)
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        flush.fill_(1)
        a.record()
        fs.sample(zp_d, zi_d, out=out_d)
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs) / iters


def run_ours(args) -> None:
    import torch
    import torch.distributed as dist
    from multimodal_diffusion_b200 import _lib
    from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig, clip_noise, sample_sharded, shard_range
    from _common import seeded_modules

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL announces itself ("NCCL version ...") on STDOUT at its first communicator: keep stdout to the one JSON line
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    lib = _lib.load()
    _lib.check(lib.avd_device_check())
    if args.chains > 0:
        _lib.set_option("chains", args.chains)

    target = "audio" if args.direction == "v2a" else "video"
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    cfg = HotPathConfig(sampler_steps=SAMPLER_STEPS, guidance=GUIDANCE)
    mods = seeded_modules()                      # random-init weights of the mvp architecture (no checkpoint exists)
    samplers = {}

    def factory(batch, clip_index0, _cfg=cfg, _target=target):
        """sample_sharded's sampler_factory: one captured FusedSampler per (config, direction, block size) on this rank."""
        key = (id(_cfg), _target, batch)
        if key not in samplers:
            samplers[key] = FusedSampler(_cfg, mods[0], mods[1], mods[2], mods[3], _target, batch, device=dev,
                                         core_precision=args.core_precision, head_precision=args.head_precision,
                                         clip_index0=clip_index0)
        return samplers[key]

    fs = factory(B, rank * B)
    # per-clip seeded synthetic latents: clip index = rank * B + b, so the result does not depend on the sharding.
    # Every rank holds the inputs of ALL world * B clips (host, pinned); sample_sharded reads only this rank's block.
    n_all = world * B
    p_shape, z_shape = fs.prompt_shape[1:], fs.latent_shape[1:]
    zp_h = torch.stack([clip_noise(p_shape, i, 1000) for i in range(n_all)]).pin_memory()
    zi_h = torch.stack([clip_noise(z_shape, i, 2000) for i in range(n_all)]).pin_memory()
    out_h = torch.empty((n_all,) + tuple(z_shape), dtype=torch.float32).pin_memory()
    zp_d, zi_d = zp_h.to(dev), zi_h.to(dev)
    out_d = torch.empty((n_all,) + tuple(z_shape), device=dev)
    flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)      # > 126 MB L2

    n0 = lib.avd_launch_count()
    fs.capture()
    launches_per_call = int(lib.avd_launch_count() - n0) - (1 + fs.kernels_per_sample() // cfg.sampler_steps)
    # (capture() runs one un-captured prepare + step for warm-up; the rest is what each graph replay launches)

    def one_call(host_io: bool):
        # the product-level clip-sharded sampler: this rank's block through its FusedSampler (H2D of the block when the
        # inputs are host tensors), then ONE all-gather of the final latents over NCCL
        sample_sharded(factory, zp_h if host_io else zp_d, zi_h if host_io else zi_d, rank=rank, world=world, out=out_d)
        if host_io:
            out_h.copy_(out_d, non_blocking=True)

    def timed(host_io: bool, iters: int):
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        for a, b in evs:
            flush.fill_(1)                                   # evict L2 between timed iterations (not timed)
            a.record()
            one_call(host_io)
            b.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(W):
        one_call(False)
        one_call(True)
    torch.cuda.synchronize()
    with ClockSampler(local) as clk:
        ms_dev = timed(False, K)
        ms_e2e = timed(True, K)
    clocks = clk.summary()
    fs.check_errors()
    value = world * B * K / (ms_dev * 1e-3)
    e2e = world * B * K / (ms_e2e * 1e-3)

    # ---- shard check (N > 1): rank 0 re-samples the LAST rank's clip block alone and compares it with the gathered slice
    shard_check = None
    if world > 1:
        one_call(False)
        torch.cuda.synchronize()
        if rank == 0:
            lo, hi = shard_range(n_all, world - 1, world)
            again = fs.sample(zp_d[lo:hi], zi_d[lo:hi])
            same = torch.equal(again, out_d[lo:hi])
            err = float((again.double() - out_d[lo:hi].double()).norm() / out_d[lo:hi].double().norm())
            shard_check = "ok" if same else (f"ok (rel-L2 {err:.1e})" if err < 1e-5 else f"MISMATCH rel-L2 {err:.3e}")
        dist.barrier()

    # ---- live per-kernel-family timing.  Kernels last 3-10 us, so an event pair per launch would mostly time the
    # event nodes.  Instead, for every kernel family the same sampling call is captured into a CUDA graph that holds
    # ONLY that family's launches (avd_sampler.profile_skip_kinds drops the others for calls on THIS sampler; same
    # launch order, arguments and programmatic-dependent-launch chain), and CUDA events on the launching stream time
    # whole replays of it: average launch duration = elapsed / launches.
    import ctypes as C
    prof = {}
    if rank == 0:
        zp_g, z_g = zp_d[:B].clone(), zi_d[:B].clone()
        n_kinds = lib.avd_prof_num_kinds()
        for kind in range(n_kinds):
            fs.args.profile_skip_kinds = ((1 << n_kinds) - 1) & ~(1 << kind)
            n_before = lib.avd_launch_count()
            torch.cuda.synchronize()
            pg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(pg):
                _lib.check(lib.avd_sample_loop(C.byref(fs.args), zp_g.data_ptr(), z_g.data_ptr(), _lib.stream_ptr()))
            n_launch = int(lib.avd_launch_count() - n_before)
            if n_launch == 0:
                continue
            pg.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 3
            e0.record()
            for _ in range(reps):
                pg.replay()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            prof[lib.avd_prof_kind_name(kind).decode()] = {"launches": n_launch, "ms": ms,
                                                           "us_per_launch": 1e3 * ms / n_launch}
            del pg
        fs.args.profile_skip_kinds = 0

    # ---- extra legs (every rank takes part in the sharded ones; rank 0 reports)
    extras = {}
    if not args.no_extras:
        # (a) BASELINE.json configs[2] AS WRITTEN: A2V, 64 clips in total, clip-sharded 64/32/16/8 per GPU: strong scaling
        n64 = 64
        lo, hi = shard_range(n64, rank, world)
        fa = FusedSampler(cfg, mods[0], mods[1], mods[2], mods[3], "video", hi - lo, device=dev,
                          core_precision=args.core_precision, head_precision=args.head_precision)
        zpa = torch.stack([clip_noise((8, 150), i, 1000) for i in range(lo, hi)]).to(dev)
        zia = torch.stack([clip_noise((8, 12, 16, 16), i, 2000) for i in range(lo, hi)]).to(dev)
        oa = torch.empty_like(zia)
        ga = torch.empty((world,) + tuple(zia.shape), device=dev) if world > 1 else None

        def a2v_call():
            fa.sample(zpa, zia, out=oa)
            if world > 1:
                dist.all_gather_into_tensor(ga.view(-1), oa.view(-1))

        for _ in range(2):
            a2v_call()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(3)]
        for a, b in ev:
            flush.fill_(1)
            a.record(); a2v_call(); b.record()
        torch.cuda.synchronize()
        ms_a = sum(a.elapsed_time(b) for a, b in ev) / 3
        if world > 1:
            t = torch.tensor([ms_a], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_a = float(t.item())
        fa.check_errors()
        extras["strong_a2v_b64"] = {"workload": "configs/a2v.yaml A2V, 64 clips TOTAL clip-sharded over the GPUs (BASELINE.json configs[2])",
                                    "clips_total": n64, "clips_per_gpu": hi - lo, "n_gpus": world, "ms_per_call": ms_a,
                                    "value": n64 / (ms_a * 1e-3), "unit": "clips/s", "scaling": "strong",
                                    "flops_per_clip": cfg.flops_per_clip("video"),
                                    "frac_of_bf16_peak": cfg.flops_per_clip("video") * n64 / (ms_a * 1e-3) / 1e12
                                                         / (world * peaks()["bf16_tflops_sustained"])}
        del fa, zpa, zia, oa, ga
        # (b) configs[3]: 10 s clip (N = 445 joint tokens), A2V, 8 clips per GPU, clip-sharded (weak)
        cfg10 = HotPathConfig(sampler_steps=SAMPLER_STEPS, guidance=GUIDANCE, Fa=500, video_latent=(40, 16, 16))
        fl = FusedSampler(cfg10, mods[0], mods[1], mods[2], mods[3], "video", 8, device=dev,
                          core_precision=args.core_precision, head_precision=args.head_precision)
        zpl = torch.stack([clip_noise((8, 500), rank * 8 + i, 1000) for i in range(8)]).to(dev)
        zil = torch.stack([clip_noise((8, 40, 16, 16), rank * 8 + i, 2000) for i in range(8)]).to(dev)
        ol = torch.empty_like(zil)
        ms_l = time_sampler_calls(fl, zpl, zil, ol, flush, 3)
        if world > 1:
            t = torch.tensor([ms_l], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_l = float(t.item())
        extras["long_context"] = {"workload": "A2V 10 s clip at 16 fps / 16 kHz: video latent [B,8,40,16,16], audio prompt [B,8,500], "
                                              "N = 445 joint tokens (BASELINE.json configs[3]), 8 clips per GPU, clip-sharded",
                                  "n_gpus": world, "ms_per_call": ms_l, "value": world * 8 / (ms_l * 1e-3), "unit": "clips/s",
                                  "flops_per_clip": cfg10.flops_per_clip("video"),
                                  "frac_of_bf16_peak": cfg10.flops_per_clip("video") * 8 / (ms_l * 1e-3) / 1e12
                                                       / peaks()["bf16_tflops_sustained"]}
        del fl, zpl, zil, ol
        # (c) eager PyTorch on the SAME GPU (rank 0): the oracle loop on device="cuda", fp32 and bf16 autocast
        if rank == 0:
            tc = {}
            for direction in ("v2a", "a2v"):
                for name, ac in (("fp32_eager", False), ("bf16_autocast", True)):
                    sub = 10
                    sec = torch_cuda_seconds(direction, B, sub, ac)
                    tc[f"{direction}_{name}"] = {"value": B / (sec * SAMPLER_STEPS / sub), "unit": "clips/s",
                                                 "ms_per_ddim_step": 1e3 * sec / sub}
            tc["what"] = (f"the CPU port's loop (sample_clip.py:317-389 restated) run UNCHANGED on this GPU with eager PyTorch, batch {B}: "
                          "cuBLAS GEMMs, F.scaled_dot_product_attention, the reference's Python un-tokenise loop nest (V2A); "
                          "10 of 50 DDIM steps timed with CUDA events, scaled x5; TF32 off for fp32_eager")
            extras["torch_cuda_baseline"] = tc
    if world > 1:
        dist.barrier()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    pk = peaks()
    total_prof_ms = ms_dev / K          # one full sampling call (graph replay, all families, L2 flushed before it)
    g = prof.get("gemm_tcgen05", {"launches": 1, "ms": float("nan")})
    gemm_tflops = gemm_flops_per_call(cfg, target, B) / (g["ms"] * 1e-3) / 1e12
    roofline = {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05, all GEMM launches of one sampling call)",
                "achieved": gemm_tflops, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                "frac": gemm_tflops / pk["bf16_tflops_sustained"],
                "peak_source": f"{pk['source']} sustained bf16 (kernel timed inside a long step)",
                "traffic": ncu_traffic_bytes(), "launches": g["launches"], "us_per_launch": g.get("us_per_launch"),
                "share_of_step": g["ms"] / total_prof_ms,
                "timing": "CUDA events around replays of a graph holding only this family's launches of one call"}
    kernels = {k: {"launches": v["launches"], "us_per_launch": round(v["us_per_launch"], 3),
                   "share": round(v["ms"] / total_prof_ms, 4)} for k, v in prof.items()}
    if "attn_tcgen05" in prof:
        kernels["attn_tcgen05"]["tflops"] = attn_flops_per_call(cfg, B) / (prof["attn_tcgen05"]["ms"] * 1e-3) / 1e12
    if "cfg_untok_ddim" in prof:
        elems = B * (cfg.Ca * cfg.Fa if target == "audio" else cfg.Cv * cfg.video_latent[0] * cfg.video_latent[1] * cfg.video_latent[2])
        kernels["cfg_untok_ddim"]["gbs"] = 16.0 * elems * cfg.sampler_steps / (prof["cfg_untok_ddim"]["ms"] * 1e-3) / 1e9

    cpu = None
    if not args.no_cpu_baseline:
        threads = host_threads()
        sub = max(1, args.cpu_sample_steps)
        cpu_sample_seconds(args.direction, B, 1, threads)               # warm-up
        sec = cpu_sample_seconds(args.direction, B, sub, threads)
        sec1 = cpu_sample_seconds(args.direction, 1, sub, threads)      # B = 1 sequential, as the reference really runs
        cpu = {"value": B / (sec * SAMPLER_STEPS / sub), "unit": "clips/s", "cores": threads, "kind": "port",
               "sample": f"{sub} of {SAMPLER_STEPS} DDIM+CFG steps for {B} clips ({sec:.2f} s), scaled x{SAMPLER_STEPS / sub:g}; "
                         "fp32 torch-CPU oracle port of sample_clip.py:317-389 (batched re-drive)",
               "b1_sequential": {"value": 1.0 / (sec1 * SAMPLER_STEPS / sub), "unit": "clips/s",
                                 "sample": f"{sub} of {SAMPLER_STEPS} steps at batch 1 ({sec1:.2f} s), scaled: the reference's own "
                                           "mode of operation (sample_clip.py:281, one clip at a time)"}}

    flops_call = cfg.flops_per_clip(target) * B
    h2d = (zp_h[:B].numel() + zi_h[:B].numel()) * 4
    d2h = out_h.numel() * 4
    line = {
        "metric": "DDIM+CFG clips/sec (mvp.yaml)", "value": value, "unit": "clips/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": ms_dev / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": args.core_precision, "data": "synthetic",
        "config": {"workload": workload_name(args.direction, B), "direction": args.direction, "batch_per_gpu": B,
                   "sampler_steps": SAMPLER_STEPS, "core_precision": args.core_precision,
                   "head_precision": args.head_precision,
                   "precision_note": "16-bit tensor-core operands (tcgen05 kind::f16: fp16 and bf16 run at the same rate), "
                                     "fp32 accumulation in TMEM, fp32 residual stream / norms / softmax / DDIM",
                   "weights": "random init (seeded), mvp.yaml architecture",
                   "l2": "flushed between timed iterations (256 MiB fill, untimed)", "cuda_graph": True,
                   "chains": _lib.get_option("chains"),
                   "api": "multimodal_diffusion_b200.sampler.sample_sharded (FusedSampler per rank + one all-gather)",
                   "parallelism": f"clip-sharded x{world}, all-gather of final latents" if world > 1 else "single GPU"},
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": "clips/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / K},
        "gpu_launches": launches_per_call * K,
        "kernels_per_ddim_step": launches_per_call // SAMPLER_STEPS,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "kernels": kernels,
        "step_tflops": flops_call / (ms_dev / K * 1e-3) / 1e12,
        "step_frac_of_bf16_peak": flops_call / (ms_dev / K * 1e-3) / 1e12 / pk["bf16_tflops_sustained"],
    }
    if shard_check is not None:
        line["shard_check"] = shard_check
    line.update(extras)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
