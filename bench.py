#!/usr/bin/env python
"""Headline benchmark: images/sec @512x512, 50-step DDIM, CFG 7.5, stablediffusionv2 512-base UNet (BASELINE.json
configs[2]: batch 64 per GPU, the configuration the metric is quoted on; it fits one B200).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the hot path over one batch: `DDIMSampler.sample(S=50, batch_size=64, ...)` — 50 guided UNet
evaluations (CFG batch 128) and 50 fused combine+update kernels.  Synthetic prompts embeddings / latents and seeded
random-init weights (no checkpoints ship with the reference).

  value : whole-job images/sec with x_T and the conditioning already resident in HBM (CUDA events, max over ranks)
  e2e   : the same through the public sampler API starting from pinned HOST buffers, H2D of x_T / c / uc and D2H
          of the latents inside the timed region
  roofline : the dominant kernel (tcgen05 implicit GEMM, mdx_gemm_tc): sum of algorithmic FLOPs over every launch
          of one UNet evaluation / sum of their CUDA-event durations, measured live right after the timed region
  cpu_baseline : the oracle (torch fp32 restatement of the reference; MindSpore is not installable here) on the host
          cores for a bounded sample of the same workload
`--impl reference` times that CPU oracle as the reference arm (rank 0 only).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PAIR_TIMEOUT = 240   # seconds the CFG pair-split phase (N >= 2) may take before every rank gives up on it
FLOP_PER_SAMPLE_FWD_SD2_64 = 804.3e9  # SURVEY.md §3.4 / BASELINE.md §3 (2*MAC), per sample per UNet forward

SD2_UNET = dict(in_channels=4, out_channels=4, model_channels=320, attention_resolutions=[4, 2, 1],
                num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_head_channels=64, num_heads=-1,
                use_spatial_transformer=True, use_linear_in_transformer=True, transformer_depth=1,
                context_dim=1024, legacy=False)  # configs/v2-inference.yaml:21-37


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    # overrides for development runs; the defaults ARE the named workload
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--sampler-steps", type=int, default=50)
    ap.add_argument("--latent", type=int, default=64)
    ap.add_argument("--scale", type=float, default=7.5)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--cfg-split", action="store_true",
                    help="N >= 2: split the cond/uncond halves of each batch across GPU pairs (per-step eps exchange)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-pair-split", action="store_true", help="N >= 2: skip the CFG pair-split measurement")
    return ap.parse_args()


def workload_name(a):
    return "stablediffusionv2 512-base UNet, batch %d/GPU, %d-step DDIM, CFG %.1f, %dx%d latent" % (
        a.batch, a.sampler_steps, a.scale, a.latent, a.latent)


# ----------------------------------------------------------------------------------------------------------
# CPU arm: the oracle on the host cores, bounded sample
# ----------------------------------------------------------------------------------------------------------
def cpu_sample_seconds(a, repeats=1):
    """One guided UNet evaluation for ONE image (CFG batch 2, latent a.latent^2) + the DDIM update, torch fp32 on
    all host threads.  Returns (seconds per sample, cores)."""
    import numpy as np
    import torch
    from minddiffusion_b200.weights import random_unet_params
    from oracle import samplers as osamp
    from oracle import sd_unet
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    P = {k: torch.from_numpy(v) for k, v in random_unet_params(SD2_UNET, 0).items()}
    rs = np.random.RandomState(42)
    x = torch.from_numpy(rs.randn(1, 4, a.latent, a.latent).astype(np.float32))
    c = torch.from_numpy(np.random.RandomState(43).randn(1, 77, 1024).astype(np.float32))
    uc = torch.from_numpy(np.random.RandomState(44).randn(1, 77, 1024).astype(np.float32))
    sched = osamp.register_schedule(1000, 0.00085, 0.012)
    fn = lambda xx, t, cc: sd_unet.unet_forward(P, SD2_UNET, xx, t.to(torch.float32), cc)
    times = []
    with torch.no_grad():
        for _ in range(repeats):
            t0 = time.perf_counter()
            # one DDIM step of the 50-step schedule == plms_sample(ddim=True) restricted to its first step
            e = osamp._cfg_model_output(fn, x, torch.full((1,), 981, dtype=torch.int64), c, uc, a.scale)
            ac = sched["alphas_cumprod"]
            a_t, a_prev = float(ac[981]), float(ac[961])
            x0 = (x - (1 - a_t) ** 0.5 * e) / a_t ** 0.5
            x = a_prev ** 0.5 * x0 + (1 - a_prev) ** 0.5 * e
            times.append(time.perf_counter() - t0)
    return times, cores


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    for _ in range(a.warmup):
        cpu_sample_seconds(a, 1)
    times, cores = cpu_sample_seconds(a, a.steps)
    per_eval = sum(times) / len(times)
    ips = 1.0 / (a.sampler_steps * per_eval)
    sample = ("1 image, 1 of %d DDIM steps per bench step (one CFG UNet evaluation at batch 2 + update), "
              "images/sec extrapolated x%d" % (a.sampler_steps, a.sampler_steps))
    line = {
        "impl": "reference", "metric": "images/sec @512^2 50-step DDIM CFG7.5", "value": ips, "unit": "images/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": per_eval * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(a), "executor": "CPU restatement (torch fp32 oracle), not MindSpore"},
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel's largest shape, from the
# `ncu --set full` capture named in the note (a profiler number, copied here; everything else in the line is live)
TRAFFIC = {"bytes_per_launch": 639.2e6,
           "note": "ncu --set full, gemm_tc_kernel conv3x3 C=320 at 64x64, CFG batch 128 (M=524288, N=320, K=2880): "
                   "launched as a CTA pair (gemm_tc_kernel<160, 0, 1>, captured before the halo boxes, which change L2 -> SM traffic, not DRAM traffic): 337.7 MB read + 301.4 MB written = 639.2 MB DRAM traffic vs 673 MB algorithmic (A once + out once; weights from L2) — no re-reads; "
                   "profiles/r02_ncu_full_gemm_tc_conv3x3_c320_cg2_raw.csv"}


def hbm_probe(torch, ops, dev, B, h, hbm_peak):
    """Achieved HBM GB/s of the memory-bound kernels at the workload's shapes: algorithmic bytes / CUDA-event time of
    single launches, L2 flushed (a 512 MB fill) before each one.  The sampler kernels' tensors are small (25 MB at
    batch 64), so inside the real step they mostly hit L2; this is the cold figure."""
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    def time_one(fn, reps=5):
        best = 1e30
        for _ in range(reps):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best

    out = {}
    n = 2 * B
    x = torch.randn(B, 4, h, h, device=dev)
    eps = torch.randn(n, h, h, 16, device=dev)
    row = ops.plms_coef_row((1., 0., 0., 0.), (0.5, 0.6, 0.0, 0.70710678))
    xo, p0 = torch.empty_like(x), torch.empty_like(x)
    nb = x.numel() * 4
    cases = [("plms_step_kernel (DDIM: eps_u, eps_c, x -> x_prev, pred_x0)", 5 * nb,
              lambda: ops.sampler_step_ex("plms", eps, x, xo, 4, True, 7.5, row=row, pred_x0=p0)),
             ("dpm_step_kernel (eps_u, eps_c, x, x0_prev -> x, x0)", 6 * nb,
              lambda: ops.sampler_step_ex("dpm", eps, x, xo, 4, True, 7.5, row=[0.9, 0.4, 1.0, 0.1, 0.05, 1.0] + [0.0] * 10,
                                          hist=(p0, None, None), pred_x0=p0))]
    C = 320
    a32 = torch.randn(n, h, h, C, device=dev)
    a16 = a32.to(torch.bfloat16)
    g, b_ = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    rows = n * h * h
    o16 = torch.empty(n, h, h, C, dtype=torch.bfloat16, device=dev)
    scratch = torch.empty(n * 64, dtype=torch.float64, device=dev)
    chan = torch.zeros((n, ops.stat_slots(n, h, h), 2, C), device=dev)
    chan[:, :, 1] = 1.0
    cases += [("gn_apply_kernel fp32->bf16 + SiLU (C=320, 64x64)", rows * C * 6,
               lambda: ops.groupnorm(a32, g, b_, 32, 1e-5, silu=True, out=o16, scratch=scratch, chan1=chan)),
              ("gn_apply_kernel bf16->bf16 + SiLU (C=320, 64x64)", rows * C * 4,
               lambda: ops.groupnorm(a16, g, b_, 32, 1e-5, silu=True, out=o16, scratch=scratch, chan1=chan)),
              ("layernorm_kernel fp32->bf16 (C=320)", rows * C * 6,
               lambda: ops.layernorm(a32.view(rows, C), g, b_, out=o16.view(rows, C)))]
    for name, nbytes, fn in cases:
        fn()
        ms = time_one(fn)
        gbs = nbytes / (ms / 1e3) / 1e9
        out[name] = {"GB/s": round(gbs, 1), "frac_of_hbm_peak": round(gbs / hbm_peak, 3), "us": round(ms * 1e3, 1),
                     "algorithmic_MB": round(nbytes / 1e6, 1)}
    out["note"] = ("gn_apply includes its gn_fold launch (statistics arrive from the producer's epilogue); "
                   "peak %.1f GB/s = MEASURED_PEAKS.json copy bandwidth" % hbm_peak)
    return out


def run_b200(a):
    import numpy as np
    import torch
    import torch.distributed as dist
    from minddiffusion_b200 import _lib, ops, samplers
    from minddiffusion_b200.ldm import LatentDiffusion
    from minddiffusion_b200.unet import UNetModel
    from minddiffusion_b200.weights import random_unet_params

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    P = random_unet_params(SD2_UNET, 0)
    unet = UNetModel(P, dtype=torch.bfloat16, device=dev, **SD2_UNET)
    del P
    unet.use_cuda_graph = not a.no_graph
    ldm = LatentDiffusion(unet, timesteps=1000, linear_start=0.00085, linear_end=0.012, scale_factor=0.18215)
    groups = world  # independent prompt batches in flight
    seed_id = rank
    if a.cfg_split:
        if world < 2 or world % 2:
            raise SystemExit("--cfg-split needs an even number of GPUs")
        from minddiffusion_b200.parallel import CfgPairSplit
        ldm.cfg_split = CfgPairSplit()
        groups, seed_id = world // 2, rank // 2  # both ranks of a pair work on the same prompt batch
    sampler = samplers.DDIMSampler(ldm)

    B, S, h = a.batch, a.sampler_steps, a.latent
    shape = (4, h, h)
    # synthetic inputs of the reference shapes: x_T seed 42 (+rank), c / uc seeds 43 / 44
    x_host = torch.from_numpy(np.random.RandomState(42 + seed_id).randn(B, *shape).astype(np.float32)).pin_memory()
    c_host = torch.from_numpy(np.random.RandomState(43).randn(B, 77, 1024).astype(np.float32)).pin_memory()
    uc_host = torch.from_numpy(np.random.RandomState(44).randn(B, 77, 1024).astype(np.float32)).pin_memory()
    out_host = torch.empty((B, *shape), dtype=torch.float32).pin_memory()
    x_dev, c_dev, uc_dev = x_host.to(dev), c_host.to(dev), uc_host.to(dev)

    def step_resident():
        return sampler.sample(S, B, shape, c_dev, x_T=x_dev, verbose=False, unconditional_guidance_scale=a.scale,
                              unconditional_conditioning=uc_dev, eta=0.0)[0]

    def step_e2e():
        xs = x_host.to(dev, non_blocking=True)
        cs = c_host.to(dev, non_blocking=True)
        us = uc_host.to(dev, non_blocking=True)
        out = sampler.sample(S, B, shape, cs, x_T=xs, verbose=False, unconditional_guidance_scale=a.scale,
                             unconditional_conditioning=us, eta=0.0)[0]
        out_host.copy_(out, non_blocking=True)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(max(a.warmup, 1)):
        step_resident()
    launches0 = _lib.launch_count() + unet.replayed_launches
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms = timed(step_resident, a.steps)
    launches = _lib.launch_count() + unet.replayed_launches - launches0
    ms_e2e = timed(step_e2e, a.steps)
    clk = clocks.stop() if rank == 0 else None

    images = B * groups * a.steps
    value = images / (ms / 1e3)
    e2e = images / (ms_e2e / 1e3)

    # ---- dominant kernel, live: every mdx_gemm_tc launch of one guided UNet evaluation ----
    prof = []
    ops.GEMM_TC_PROFILE = prof
    unet.use_cuda_graph = False
    rep = 1 if a.cfg_split else 2
    t_probe = torch.full((rep * B,), 501.0, device=dev)
    if rep == 2:
        unet.set_context(torch.cat((uc_dev, c_dev), 0))   # the pair-split phase left a single-branch context behind
    unet.forward_nhwc(x_dev, t_probe, repeat=rep)
    torch.cuda.synchronize()
    ops.GEMM_TC_PROFILE = None
    flops = sum(p_[0] for p_ in prof)
    gemm_ms = sum(p_[1].elapsed_time(p_[2]) for p_ in prof)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
    peak_burst = float(peaks.get("bf16_tflops", 1650.0))
    hbm_peak = float(peaks.get("hbm_gbs", 6500.0))
    peak_kind = "measured sustained (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md, sustained)"
    achieved = flops / (gemm_ms / 1e3) / 1e12 if gemm_ms > 0 else 0.0
    hbm_kernels = hbm_probe(torch, ops, dev, B, h, hbm_peak) if world == 1 else None
    fwd_flops = FLOP_PER_SAMPLE_FWD_SD2_64 * (h / 64.0) ** 2  # conv/linear scale ~ pixels; attention differs slightly
    step_tflops = value / world * S * 2 * FLOP_PER_SAMPLE_FWD_SD2_64 / 1e12 if h == 64 else None  # per GPU

    # ---- the line as measured so far (rank 0); the pair-split phase below can only add to it ----
    base_line = None
    if rank == 0:
        cpu = None
        if not a.no_cpu_baseline and world == 1:
            times, cores = cpu_sample_seconds(a, 1)
            cpu = {"value": 1.0 / (S * times[0]), "unit": "images/s", "cores": cores, "kind": "port",
                   "sample": "one CFG UNet evaluation (batch 2, %dx%d latent) + DDIM update for 1 image = 1 of %d steps, "
                             "extrapolated x%d; oracle (torch fp32 restatement), MindSpore not installable" % (h, h, S, S)}

        base_line = {
            "metric": "images/sec @512^2 50-step DDIM CFG7.5", "value": value, "unit": "images/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(a), "global_batch": B * groups, "sampler": "DDIM", "sampler_steps": S,
                       "cfg_scale": a.scale, "parallelism": ("dp%d x cfg-pair-split (per-step 2-rank eps all-gather over NCCL)" % groups) if a.cfg_split
                       else "dp%d (independent prompt batches, no collective)" % world,
                       "l2": "inputs larger than L2 (weights 1.7 GB + activations >> 126 MB)",
                       "cuda_graph": not a.no_graph},
            "clocks": clk,
            "e2e": {"value": e2e, "unit": "images/s",
                    "h2d_bytes_per_step": (x_host.numel() + c_host.numel() + uc_host.numel()) * 4,
                    "d2h_bytes_per_step": out_host.numel() * 4},
            "gpu_launches": int(launches),
            "unet_eval_ms": ms / a.steps / S,
            "unet_tflops_per_gpu": step_tflops,
            "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel (mdx_gemm_tc: conv3x3 / 1x1 / dense)",
                         "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
                         "peak_kind": peak_kind, "frac_of_burst_peak": achieved / peak_burst, "peak_burst": peak_burst,
                         "traffic": TRAFFIC["bytes_per_launch"], "traffic_note": TRAFFIC["note"],
                         "launches": len(prof), "kernel_ms_per_unet_eval": gemm_ms,
                         "whole_step_frac": (step_tflops / peak) if step_tflops else None,
                         "whole_step_frac_of_burst_peak": (step_tflops / peak_burst) if step_tflops else None},
            "hbm_kernels": hbm_kernels,
            "cpu_baseline": cpu,
        }

    # ---- N >= 2, same run: CFG pair split (BASELINE config #3).  GPU 2k evaluates the unconditional half of a batch
    #      of B prompts, GPU 2k+1 the conditional half; each step the 4-channel eps halves are exchanged with one NCCL
    #      all-gather issued INSIDE the captured step graph and both ranks apply the same update. ----
    pair = None
    watchdog = None
    if world >= 2 and world % 2 == 0 and not a.cfg_split and not a.no_pair_split:
        from minddiffusion_b200.parallel import CfgPairSplit
        # Everything above is measured; what follows adds collectives.  A rank that never comes back from one must not
        # cost the run its line: after PAIR_TIMEOUT seconds every rank leaves, rank 0 printing the line it has.
        def give_up():
            if rank == 0:
                base_line["pair_split"] = {"error": "pair-split phase did not finish within %d s" % PAIR_TIMEOUT}
                print(json.dumps(base_line), flush=True)
            os._exit(0)
        watchdog = threading.Timer(PAIR_TIMEOUT, give_up)
        watchdog.daemon = True
        watchdog.start()
        unet.use_cuda_graph = not a.no_graph
        ldm.cfg_split = split = CfgPairSplit()
        xp = torch.from_numpy(np.random.RandomState(42 + rank // 2).randn(B, *shape).astype(np.float32)).to(dev)
        step_pair = lambda: sampler.sample(S, B, shape, c_dev, x_T=xp, verbose=False, eta=0.0,
                                           unconditional_guidance_scale=a.scale, unconditional_conditioning=uc_dev)[0]
        lat = step_pair()                                       # warm-up: captures the pair-split step graph
        both = [torch.empty_like(lat) for _ in range(2)]
        dist.all_gather(both, lat, group=split.group)
        identical = bool(torch.equal(both[0], both[1]))        # the two ranks of a pair hold the same latents
        kp = max(1, min(a.steps, 3))
        ms_pair = timed(step_pair, kp)
        probe = torch.randn(B, h, h, 16, device=dev)
        for _ in range(3):
            split.exchange(probe, channels=4)
        barrier()
        x0e, x1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        x0e.record()
        for _ in range(20):
            split.exchange(probe, channels=4)
        x1e.record()
        barrier()
        pair = {"value": B * (world // 2) * kp / (ms_pair / 1e3), "unit": "images/s", "steps": kp,
                "ms_per_unet_eval": ms_pair / kp / S, "dp_ms_per_unet_eval": ms / a.steps / S,
                "exchange_us": x0e.elapsed_time(x1e) * 1e3 / 20, "exchange_bytes_per_rank": B * h * h * 4 * 4,
                "latents_identical_on_pair": identical, "cuda_graph": not a.no_graph,
                "exchange_in_graph": bool(split.in_graph),
                "parallelism": "dp%d x cfg-pair-split (packed 4-channel eps all-gather over NCCL, %s)"
                               % (world // 2, "captured inside the step graph" if split.in_graph
                                  else "issued between the two captured segments of a step")}
        ldm.cfg_split = None
        watchdog.cancel()

    if rank == 0:
        if pair is not None:
            base_line["pair_split"] = pair
        print(json.dumps(base_line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
