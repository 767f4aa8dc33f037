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
    ap.add_argument("--cpu-sample-steps", type=int, default=4, help="DDIM steps of the CPU-baseline sample")
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
    """--impl reference: the reference's CPU algorithm (oracle port, all host threads) on a bounded sample of the
    same workload: each timed step = `cpu_sample_steps` of the 50 DDIM steps for the full batch, scaled to clips/s."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    sub = max(1, min(SAMPLER_STEPS, args.cpu_sample_steps // 2 or 1))
    for _ in range(args.warmup):
        cpu_sample_seconds(args.direction, args.batch, 1, threads)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_sample_seconds(args.direction, args.batch, sub, threads)
    per_call = t / args.steps * (SAMPLER_STEPS / sub)          # seconds for one full 50-step sampling call
    value = args.batch / per_call
    sample = (f"{sub} of {SAMPLER_STEPS} DDIM+CFG steps for the full batch of {args.batch} clips per timed step, "
              f"scaled x{SAMPLER_STEPS / sub:g}; fp32 torch-CPU oracle port of sample_clip.py:317-389")
    line = {"metric": "DDIM+CFG clips/sec (mvp.yaml)", "value": value, "unit": "clips/s", "impl": "reference",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_call * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.direction, args.batch), "direction": args.direction,
                       "batch_per_gpu": args.batch, "sampler_steps": SAMPLER_STEPS},
            "cpu_baseline": {"value": value, "unit": "clips/s", "cores": threads, "kind": "port", "sample": sample},
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


def run_ours(args) -> None:
    import torch
    import torch.distributed as dist
    from multimodal_diffusion_b200 import _lib
    from multimodal_diffusion_b200.sampler import FusedSampler, HotPathConfig, clip_noise
    from _common import seeded_modules

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    _lib.check(lib.avd_device_check())

    target = "audio" if args.direction == "v2a" else "video"
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    cfg = HotPathConfig(sampler_steps=SAMPLER_STEPS, guidance=GUIDANCE)
    mods = seeded_modules()                      # random-init weights of the mvp architecture (no checkpoint exists)
    fs = FusedSampler(cfg, mods[0], mods[1], mods[2], mods[3], target, B, device=dev,
                      core_precision=args.core_precision, head_precision=args.head_precision)
    # per-clip seeded synthetic latents: clip index = rank * B + b, so the result does not depend on the sharding
    p_shape, z_shape = fs.prompt_shape[1:], fs.latent_shape[1:]
    zp_h = torch.stack([clip_noise(p_shape, rank * B + b, 1000) for b in range(B)]).pin_memory()
    zi_h = torch.stack([clip_noise(z_shape, rank * B + b, 2000) for b in range(B)]).pin_memory()
    out_h = torch.empty(fs.latent_shape, dtype=torch.float32).pin_memory()
    zp_d, zi_d = zp_h.to(dev), zi_h.to(dev)
    out_d = torch.empty(fs.latent_shape, device=dev)
    gathered = torch.empty((world,) + tuple(fs.latent_shape), device=dev) if world > 1 else None
    flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)      # > 126 MB L2

    n0 = lib.avd_launch_count()
    fs.capture()
    launches_per_call = int(lib.avd_launch_count() - n0) - (1 + fs.kernels_per_sample() // cfg.sampler_steps)
    # (capture() runs one un-captured prepare + step for warm-up; the rest is what each graph replay launches)

    def one_call(host_io: bool):
        if host_io:
            fs.sample(zp_h, zi_h, out=out_d)
        else:
            fs.sample(zp_d, zi_d, out=out_d)
        if world > 1:
            dist.all_gather_into_tensor(gathered.view(-1), out_d.view(-1))
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
    value = world * B * K / (ms_dev * 1e-3)
    e2e = world * B * K / (ms_e2e * 1e-3)

    # ---- live per-kernel-family timing.  Kernels last 3-10 us, so an event pair per launch would mostly time the
    # event nodes.  Instead, for every kernel family the same sampling call is captured into a CUDA graph that holds
    # ONLY that family's launches (avd_debug_skip_kinds drops the others; same launch order, arguments and
    # programmatic-dependent-launch chain), and CUDA events on the launching stream time whole replays of it:
    # average launch duration = elapsed / launches.  The families' sum is compared with the full call below.
    import ctypes as C
    prof = {}
    if rank == 0:
        zp_g, z_g = zp_d.clone(), zi_d.clone()
        n_kinds = lib.avd_prof_num_kinds()
        for kind in range(n_kinds):
            _lib.check(lib.avd_debug_skip_kinds(((1 << n_kinds) - 1) & ~(1 << kind)))
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
        _lib.check(lib.avd_debug_skip_kinds(0))
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
        cpu = {"value": B / (sec * SAMPLER_STEPS / sub), "unit": "clips/s", "cores": threads, "kind": "port",
               "sample": f"{sub} of {SAMPLER_STEPS} DDIM+CFG steps for {B} clips ({sec:.2f} s), scaled x{SAMPLER_STEPS / sub:g}; "
                         "fp32 torch-CPU oracle port of sample_clip.py:317-389"}

    flops_call = cfg.flops_per_clip(target) * B
    h2d = (zp_h.numel() + zi_h.numel()) * 4
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
                   "parallelism": f"clip-sharded x{world}, all-gather of final latents" if world > 1 else "single GPU"},
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": "clips/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / K},
        "gpu_launches": launches_per_call * K,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "kernels": kernels,
        "step_tflops": flops_call / (ms_dev / K * 1e-3) / 1e12,
        "step_frac_of_bf16_peak": flops_call / (ms_dev / K * 1e-3) / 1e12 / pk["bf16_tflops_sustained"],
    }
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
