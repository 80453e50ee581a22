"""bench.py — fwd+bwd of the differentiable Gaussian-splatting rasterizer on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --steps K --warmup W    # the CPU arm (oracle port, host cores)

Workload (BASELINE.json configs[2], the one the metric is quoted on): 3 M synthetic Gaussians,
SH degree 3, 1920x1080, one view per step per GPU, forward + backward with a fixed seeded
dL/dpixel.  A "step" is one fwd+bwd of one view on every rank; with N > 1 each rank renders its
own camera of the batch (weak scaling, one camera subset per GPU) and the 59-float-per-Gaussian
parameter gradients are summed with ONE NCCL all-reduce inside the timed region.
Prints ONE JSON line on rank 0 (contract: see DESIGN.md "Measurement").
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from mvs_gaussian_splatting_b200 import synthetic as S  # noqa: E402  (numpy only)

METRIC = "rasterizer_fwd_bwd_views_per_s"
UNIT = "views/s"

# Algorithmic cost model (DESIGN.md "Kernels and rooflines").
FLOP_PER_EVAL = 18.0       # quadratic form, exp, opacity mul, min, compares, 1-alpha, T*
FLOP_PER_BLEND_FWD = 8.0   # weight + 3 fma
FLOP_PER_BLEND_BWD = 73.0  # reference backward per contributing pair (counted in DESIGN.md)
FP32_LANES_PER_SM = 128
N_SM = 148


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--points", type=int, default=3_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--sh-degree", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the raw_entry / optimizer_step side measurements (profiler passes)")
    ap.add_argument("--forward-only", action="store_true",
                    help="render.py path (BASELINE.json configs[4]): forward only, frames/s; combine with "
                         "--points 6000000 --width 3840 --height 2160 [--sh-degree 0]")
    ap.add_argument("--multi", default="sharded", choices=["sharded", "sharded-nccl", "allreduce"],
                    help="N > 1: 'sharded' = Gaussian-sharded step (all-to-all of 100 B/Gaussian), "
                         "'allreduce' = replicated parameters + one 236 B/Gaussian gradient all-reduce")
    ap.add_argument("--cpu-budget-s", type=float, default=200.0,
                    help="--impl reference: wall-clock budget the K+W CPU steps are sized for")
    return ap.parse_args()


def bench_camera(k: int, W: int, H: int) -> S.Camera:
    """Camera k of the batch: k = 0 is the front camera the scene was generated for; the others sit
    on a 0.5-unit ring around it and look at the scene centre (similar work per rank)."""
    if k == 0:
        return S.front_camera(W, H)
    a = 2 * math.pi * (k - 1) / 7.0
    R, t = S.look_at([0.5 * math.cos(a), 0.5 * math.sin(a), 0.0], [0.0, 0.0, 10.5])
    return S.make_camera(R, t, math.radians(60.0), W, H)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm_gbs=float(d["hbm_gbs"]), sm_max_mhz=float(d.get("sm_max_mhz", 1965.0)), source="measured")
    return dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, source="fallback")


# ------------------------------------------------------------------------------------------
# clocks during the timed region (NVML; same fields as the profiling recipe's nvidia-smi line)
# ------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index: int, period_s: float = 0.004):
        super().__init__(daemon=True)
        self.index, self.period, self.samples, self.reasons = index, period_s, [], set()
        self.stop_flag = threading.Event()
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self.stop_flag.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self.stop_flag.wait(self.period)

    def result(self):
        self.stop_flag.set()
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle port (oracle/gs_oracle.c) on the host cores.  The reference's own CPU/CUDA
# implementation of this path is an un-vendored submodule, so kind = "port".
# ------------------------------------------------------------------------------------------
def host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


class CpuArm:
    """fwd+bwd of one view with the oracle; the two blend stages may run on a strided subset of the
    tile rows (bounded sample), the per-Gaussian and binning stages always run in full."""

    def __init__(self, g, cam, D, dL, threads):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import gs_oracle as O
        O.build()
        self.O, self.g, self.cam, self.D, self.dL, self.threads = O, g, cam, D, dL, threads
        self.gx, self.gy = O.grid(cam.width, cam.height)

    def rows(self, stride):
        return [(r * self.gx, (r + 1) * self.gx) for r in range(0, self.gy, stride)]

    def step(self, stride=1):
        O, g, cam, th = self.O, self.g, self.cam, self.threads
        W, H = cam.width, cam.height
        common = (cam.world_view_transform, cam.full_proj_transform, cam.camera_center, W, H, cam.tanfovx,
                  cam.tanfovy, self.D)
        bg = np.zeros(3, np.float32)
        t0 = time.perf_counter()
        pre = O.preprocess(g["means3D"], g["scales"], g["rotations"], g["opacities"], g["shs"], *common, threads=th)
        binned = O.bin_instances(pre, W, H)
        t1 = time.perf_counter()
        tr = None if stride == 1 else self.rows(stride)
        fwd = O.blend_forward(pre, binned, bg, W, H, threads=th, tile_ranges=tr)
        gb = O.blend_backward(pre, binned, fwd, bg, self.dL, W, H, threads=th, tile_ranges=tr)
        t2 = time.perf_counter()
        O.preprocess_backward(g["means3D"], g["scales"], g["rotations"], g["shs"], *common, pre, gb, threads=th)
        t3 = time.perf_counter()
        frac = 1.0 if stride == 1 else len(tr) / self.gy
        t_full = (t1 - t0) + (t3 - t2) + (t2 - t1) / frac
        return dict(t_view_s=t_full, t_wall_s=t3 - t0, rows_frac=frac, t_fixed_s=(t1 - t0) + (t3 - t2),
                    t_blend_s=t2 - t1)

    def describe(self, stride):
        if stride == 1:
            return "full view: all Gaussians, all tiles"
        return (f"all Gaussians through preprocess/binning/preprocess-backward; blend fwd+bwd on every "
                f"{stride}th tile row ({len(self.rows(stride))}/{self.gy} rows), blend time scaled by rows")


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # under torchrun only rank 0 runs the CPU arm
    W, H, P, D = args.width, args.height, args.points, args.sh_degree
    cam = bench_camera(0, W, H)
    g = S.make_gaussians(P, cam, seed=0)
    dL = np.random.Generator(np.random.PCG64(1234)).normal(size=(3, H, W)).astype(np.float32)
    th = host_threads()
    arm = CpuArm(g, cam, D, dL, th)
    # size the per-step sample: one full calibration step (untimed), then pick the row stride
    cal = arm.step(1)
    n = args.steps + args.warmup
    per_step = max(args.cpu_budget_s - cal["t_wall_s"], 1.0) / max(n, 1)
    stride = 1
    while stride < arm.gy and arm_estimate(cal, stride, arm.gy) > per_step:
        stride += 1
    for _ in range(args.warmup):
        arm.step(stride)
    ts = [arm.step(stride)["t_view_s"] for _ in range(args.steps)]
    t = float(np.mean(ts)) if ts else cal["t_view_s"]
    value = 1.0 / t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{P} Gaussians SH{D} {W}x{H} fwd+bwd, 1 view", "arm": "CPU oracle port, host cores"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": th, "kind": "port", "sample": arm.describe(stride)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


def arm_estimate(cal, stride, gy):
    """Predicted wall seconds of one sampled step from the full calibration step."""
    rows = len(range(0, gy, stride))
    return cal["t_fixed_s"] + cal["t_blend_s"] * rows / gy


# ------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------
def stage_rooflines(stage_ms, counts, peaks, clocks):
    """Per-stage algorithmic cost / measured time against the roofline that bounds the stage.
    Bytes and FLOPs per unit are the figures of SURVEY 8(d), restated in DESIGN.md."""
    P, Pv, R, tiles = counts["P"], counts["P_visible"], counts["R"], counts["tiles"]
    E, C, En = counts["pairs_evaluated"], counts["pairs_blended"], counts["sum_n_contrib"]
    keybits = 32 + max(1, math.ceil(math.log2(max(tiles, 2))))
    hbm = {
        "preprocess": 52.0 * P + 259.0 * Pv,
        "depth_sort": 68.0 * P,          # one histogram read + 4 passes x (8 B read + 8 B write)
        "scan": 12.0 * P,                # order + gathered tiles_touched read, offsets written
        "emit_keys": 12.0 * R + 24.0 * Pv,
        # tile-bit sort only (the depth order comes from depth_sort): histogram read + passes x (12 r + 12 w)
        "radix_sort": (8.0 + 24.0 * math.ceil((keybits - 32) / 8)) * R,
        "tile_ranges": 8.0 * R + 8.0 * tiles,
        "preprocess_backward": 303.0 * Pv + 240.0 * P,
    }
    flops = {
        "blend_forward": FLOP_PER_EVAL * E + FLOP_PER_BLEND_FWD * C,
        "blend_backward": FLOP_PER_EVAL * En + FLOP_PER_BLEND_BWD * C,
    }
    fp32_peak = N_SM * FP32_LANES_PER_SM * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12  # TFLOP/s, FMA = 2
    out = {}
    for name, ms in stage_ms.items():
        if ms <= 0:
            continue
        if name in hbm:
            a = hbm[name] / (ms * 1e-3) / 1e9
            out[name] = {"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": a / peaks["hbm_gbs"], "traffic": None, "ms": ms,
                         "algorithmic_bytes": hbm[name], "peak_source": peaks["source"]}
        elif name in flops:
            a = flops[name] / (ms * 1e-3) / 1e12
            out[name] = {"bound": "fp32_simt_issue", "achieved": a, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": a / fp32_peak, "traffic": None, "ms": ms, "algorithmic_flops": flops[name],
                         "peak_source": "nominal 148 SM x 128 lanes x 2 x clocks.max.sm (no tensor-core work on "
                                        "this path; MEASURED_PEAKS.json has no FP32 SIMT figure)"}
    return out


def load_traffic():
    """dram bytes per launch from the committed ncu --set full capture, if one is recorded."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


def run_b200_arm(args):
    import torch
    import torch.distributed as dist
    from mvs_gaussian_splatting_b200 import (GaussianRasterizationSettings, GaussianRasterizer, GradientArena,
                                             _native, blend_pair_counts, last_forward_stats, rasterize_backward,
                                             rasterize_forward)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the rasterizer has no CPU path; use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    _native.lib()  # fail loudly when libgsrast_b200.so is missing

    W, H, P, D = args.width, args.height, args.points, args.sh_degree
    M = 16
    g = S.make_gaussians(P, bench_camera(0, W, H), sh_coeffs=M, seed=0)   # same scene on every rank
    cam = bench_camera(rank, W, H)                                        # this rank's camera of the batch
    host = {k: torch.from_numpy(v).pin_memory() for k, v in g.items()}
    t = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
    kw = S.settings_kwargs(cam, [0.0, 0.0, 0.0], D)
    cam_host = torch.from_numpy(np.concatenate([kw["viewmatrix"].ravel(), kw["projmatrix"].ravel(),
                                                kw["campos"].ravel(), kw["bg"].ravel()]).astype(np.float32)).pin_memory()

    def settings_from(block):
        k2 = dict(kw)
        k2.update(viewmatrix=block[0:16].view(4, 4), projmatrix=block[16:32].view(4, 4), campos=block[32:35],
                  bg=block[35:38])
        return GaussianRasterizationSettings(**k2)

    rs = settings_from(cam_host.to(dev))
    rng = np.random.Generator(np.random.PCG64(1234))
    dL_np = rng.normal(size=(3, H, W)).astype(np.float32)
    dL = torch.from_numpy(dL_np).to(dev)
    arena = GradientArena(P, M, dev)
    sharded = world > 1 and args.multi.startswith("sharded")
    sr = cams_all = slices = None
    if sharded:
        from mvs_gaussian_splatting_b200.sharded import ShardedViewRasterizer
        sr = ShardedViewRasterizer(P, M, W, H, dev, exchange="nccl" if args.multi == "sharded-nccl" else "peer")
        rows = sr.plan.rows(rank)
        slices = {k: t[k][rows].contiguous() for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        cams_all = []
        for v in range(world):
            kv = S.settings_kwargs(bench_camera(v, W, H), [0.0, 0.0, 0.0], D)
            for k in ("bg", "viewmatrix", "projmatrix", "campos"):
                kv[k] = torch.from_numpy(np.ascontiguousarray(kv[k])).to(dev)
            cams_all.append(GaussianRasterizationSettings(**kv))
        blocks = []
        for v in range(world):
            kv = S.settings_kwargs(bench_camera(v, W, H), [0.0, 0.0, 0.0], D)
            blocks += [kv["viewmatrix"].ravel(), kv["projmatrix"].ravel(), kv["campos"].ravel(), kv["bg"].ravel()]
        cam_host_all = torch.from_numpy(np.concatenate(blocks).astype(np.float32)).pin_memory()

    def step():
        if sharded:
            sr.forward(slices, cams_all)
            sr.backward(dL)
            return sr._state[0]
        color, radii, ctx, inp = rasterize_forward(t["means3D"], t["shs"], None, t["opacities"], t["scales"],
                                                   t["rotations"], None, rs)
        rasterize_backward(ctx, inp, rs, dL, arena=arena)
        if world > 1:
            dist.all_reduce(arena.flat)
        return ctx

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        ctx = step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    _native.stage_timing_enable(True)
    _native.stage_timing_read()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = _native.launch_count()
    e0.record()
    for _ in range(args.steps):
        ctx = step()
    e1.record()
    barrier()
    launches = _native.launch_count() - n0
    ms_total = e0.elapsed_time(e1)
    _native.stage_timing_enable(False)
    stages = _native.stage_timing_read()
    clocks = sampler.result()
    if world > 1:
        tt = torch.tensor([ms_total], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
    ms_step = ms_total / args.steps
    value = world * 1e3 / ms_step

    # ---- counts for the algorithmic cost model (outside the timed region) ----
    st = last_forward_stats(dev)
    ev, bl, ev_ours, warp_iters = blend_pair_counts(ctx)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    if sharded:
        st["num_visible"] = int((ctx.radii > 0).sum().item())
    counts = dict(P=P, P_visible=st["num_visible"], R=st["num_rendered"], tiles=gx * gy, pairs_evaluated=ev,
                  pairs_blended=bl, pairs_evaluated_after_culling=ev_ours, warp_splat_iterations=warp_iters,
                  sum_n_contrib=int(ctx.array("n_contrib").long().sum().item()))
    stage_ms = {k: v[0] / max(args.steps, 1) for k, v in stages.items() if v[1]}   # per step (a stage may launch N times)
    peaks = load_peaks()
    roofs = stage_rooflines(stage_ms, counts, peaks, clocks)
    traffic = load_traffic()
    for k, r in roofs.items():
        r["traffic"] = traffic.get(k)
    dominant = max(stage_ms, key=stage_ms.get)
    per_rank = None
    if world > 1:   # which rank sets the pace: each rank renders a different view (different R, different blend cost)
        mine = dict(rank=rank, num_rendered=int(counts["R"]), stage_ms_sum=round(sum(stage_ms.values()), 4),
                    blend_ms=round(stage_ms.get("blend_forward", 0.0) + stage_ms.get("blend_backward", 0.0), 4))
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        per_rank = gathered

    # ---- e2e: the public autograd API with host-side per-step inputs ----
    e2e = run_e2e(args, torch, dist, world, dev, t, cam_host, settings_from, dL_np, GaussianRasterizer,
                  sharded=(sr, slices, rank, cam_host_all) if sharded else None)
    e2e_hp = run_e2e_host_params(args, torch, world, dev, host, rs, dL, rasterize_forward, rasterize_backward,
                                 arena) if world == 1 else None

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{P} Gaussians SH{D} {W}x{H} fwd+bwd, 1 view per GPU per step"
                               + (f", Gaussian-sharded ({sr.exchange} exchange): all-to-all of 52 B (fwd) + 48 B (bwd) per Gaussian and view, "
                                  "gradients of each rank's parameter slice summed over the views" if sharded else
                                  ", + NCCL all-reduce of 59 f32/Gaussian gradients" if world > 1 else ""),
                   "baseline_config": "BASELINE.json configs[2]", "l2": "inputs (708 MB) larger than L2, no flush",
                   "parallelism": f"views x{world}", "counts": counts},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofs.get(dominant),
        "dominant_stage": dominant,
        "stages": roofs,
        "stage_ms_sum": sum(stage_ms.values()),
        "e2e": e2e,
    }
    if e2e_hp is not None:
        line["e2e_host_params"] = e2e_hp
    if per_rank is not None:
        line["per_rank"] = per_rank
    if world == 1 and not args.no_extras:
        line["raw_entry"] = run_raw_compare(args, torch, dev, t, rs, dL, GaussianRasterizer)
        line["optimizer_step"] = run_optimizer_compare(torch, dev, t)
        line["loss_step"] = run_loss_compare(torch, dev, W, H)
        line["ply_io"] = run_ply_io(torch, dev, t)
        line["train_iteration"] = run_train_iteration_compare(torch, dev, t, rs, W, H, GaussianRasterizer)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(g, bench_camera(0, W, H), D, dL_np)
    if rank == 0:
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline(g, cam, D, dL_np, budget_s=30.0):
    th = host_threads()
    arm = CpuArm(g, cam, D, dL_np, th)
    cal = arm.step(1)                      # warm-up + calibration (untimed)
    stride = 1
    while stride < arm.gy and 2 * arm_estimate(cal, stride, arm.gy) > budget_s:
        stride += 1
    ts = [arm.step(stride)["t_view_s"] for _ in range(2)]
    t = float(np.mean(ts))
    return {"value": 1.0 / t, "unit": UNIT, "cores": th, "kind": "port", "sample": arm.describe(stride),
            "ms_per_view": 1e3 * t}


def run_e2e(args, torch, dist, world, dev, t, cam_host, settings_from, dL_np, GaussianRasterizer, sharded=None):
    """The call a user of the reference makes (gaussian_renderer/__init__.py:257-265 + train.py:95-113):
    per step the camera block and the ground-truth image come from PINNED HOST memory, the image is
    rendered through GaussianRasterizer (autograd), an L1 loss is formed, backward() runs, and the
    loss is read back to the host.  The Gaussian parameters are model state and stay in HBM, as in
    the reference's training loop."""
    from mvs_gaussian_splatting_b200.loss import l1_loss
    H, W = args.height, args.width
    gt_host = torch.from_numpy((0.5 + 0.25 * dL_np).clip(0, 1).astype(np.float32)).pin_memory()
    params = {k: t[k].detach().clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    if sharded is not None:
        # Gaussian-sharded step: the host block carries the cameras of ALL N views of the step (38 floats each)
        sr, slices, my_rank, cam_host = sharded
        params = None

    # Per-step host inputs are prefetched one step ahead on a copy stream (double-buffered), as a data
    # loader does; the loss is copied to pinned memory asynchronously and read one step late, so the
    # host never stalls the launch queue.  Every copy still happens inside the timed region, once per step.
    copy_stream = torch.cuda.Stream(device=dev)
    main = torch.cuda.current_stream(dev)
    slots = [dict(block=torch.empty_like(cam_host, device=dev), gt=torch.empty_like(gt_host, device=dev),
                  ready=torch.cuda.Event(), free=torch.cuda.Event(), loss=torch.zeros(1).pin_memory(),
                  loss_ready=torch.cuda.Event()) for _ in range(2)]
    for sl in slots:
        sl["free"].record(main)
    state = {"k": 0, "pending": None}

    def prefetch(sl):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(sl["free"])          # the step that last used this slot is done with it
            sl["block"].copy_(cam_host, non_blocking=True)
            sl["gt"].copy_(gt_host, non_blocking=True)
            sl["ready"].record(copy_stream)

    prefetch(slots[0])

    def step():
        sl = slots[state["k"] & 1]
        prefetch(slots[(state["k"] + 1) & 1])           # next step's inputs travel while this step computes
        main.wait_event(sl["ready"])
        if sharded is not None:
            cams = [settings_from(sl["block"][38 * v:38 * v + 38]) for v in range(world)]
            color, radii = sr.forward(slices, cams)
            color.requires_grad_(True)
            loss = l1_loss(color, sl["gt"])                   # library loss (loss.py): one launch each way
            loss.backward()
            sr.backward(color.grad)
        else:
            rs = settings_from(sl["block"])
            means2D = torch.zeros_like(params["means3D"], requires_grad=True)
            color, radii = GaussianRasterizer(rs)(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                                  opacities=params["opacities"], scales=params["scales"],
                                                  rotations=params["rotations"])
            loss = l1_loss(color, sl["gt"])
            loss.backward()
            if world > 1:
                for p in params.values():
                    dist.all_reduce(p.grad)
        sl["loss"].copy_(loss.detach().reshape(1), non_blocking=True)   # D2H read of the step's result
        sl["loss_ready"].record(main)
        sl["free"].record(main)
        val = None
        if state["pending"] is not None:                 # previous step's loss, now on the host
            state["pending"]["loss_ready"].synchronize()
            val = float(state["pending"]["loss"][0])
        state["pending"] = sl
        state["k"] += 1
        if params is not None:
            for p in params.values():
                p.grad = None                               # train.py:142 zero_grad(set_to_none=True)
        return val

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    n = max(30, args.steps)   # wall-clock: enough steps that one host hiccup does not move the figure
    prof = None
    if os.environ.get("GSR_E2E_PROFILE"):   # where the host time of an e2e step goes (stderr; perturbs the figure)
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if prof is not None:
        import pstats
        prof.disable()
        pstats.Stats(prof, stream=sys.stderr).sort_stats("tottime").print_stats(14)
    if world > 1:
        tt = torch.tensor([dt], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
    return {"value": world * n / dt, "unit": UNIT, "ms_per_step": 1e3 * dt / n, "steps": n,
            "h2d_bytes_per_step": int(cam_host.numel() * 4 + gt_host.numel() * 4), "d2h_bytes_per_step": 4,
            "api": "GaussianRasterizer(settings)(...) + l1_loss (mvs_gaussian_splatting_b200.loss) + backward(), camera + ground-truth image from "
                   "pinned host memory (prefetched one step ahead on a copy stream), loss copied back to pinned "
                   "memory every step; Gaussian parameters resident in HBM (model state)"
                   + ("; N > 1: ShardedViewRasterizer.forward/backward on each rank's parameter slice" if sharded else "")}


def run_raw_compare(args, torch, dev, t, rs, dL, GaussianRasterizer):
    """the step as the reference's render() forms it — PyTorch getters (exp, sigmoid, normalize,
    cat of _features_dc | _features_rest; scene/gaussian_model.py:152-183) around the rasterizer, autograd
    through them — against `forward_raw`, which fuses those ops and their backward into the kernels.  Same raw
    leaf tensors, same dL/dpixel; CUDA events, 10 steps each."""
    sh = t["shs"]
    raw = dict(means3D=t["means3D"], features_dc=sh[:, :1].contiguous(), features_rest=sh[:, 1:].contiguous(),
               opacity_raw=torch.logit(t["opacities"].clamp(1e-6, 1 - 1e-6)), scaling_raw=torch.log(t["scales"]),
               rotation_raw=t["rotations"] * 1.7)
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in raw.items()}
    R = GaussianRasterizer(rs)

    def torch_step():
        m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
        shs = torch.cat((leaves["features_dc"], leaves["features_rest"]), dim=1)
        color, _ = R(means3D=leaves["means3D"], means2D=m2, shs=shs, opacities=torch.sigmoid(leaves["opacity_raw"]),
                     scales=torch.exp(leaves["scaling_raw"]),
                     rotations=torch.nn.functional.normalize(leaves["rotation_raw"]))
        color.backward(dL)

    def fused_step():
        m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
        color, _ = R.forward_raw(leaves["means3D"], m2, leaves["features_dc"], leaves["features_rest"],
                                 leaves["opacity_raw"], leaves["scaling_raw"], leaves["rotation_raw"])
        color.backward(dL)

    out = {}
    for name, fn in (("torch_activations_ms", torch_step), ("fused_raw_ms", fused_step)):
        for _ in range(3):
            fn()
            for p in leaves.values():
                p.grad = None
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
            for p in leaves.values():
                p.grad = None
        e1.record()
        torch.cuda.synchronize()
        out[name] = e0.elapsed_time(e1) / 10
    out["note"] = ("fwd+bwd of one view from raw leaf parameters: PyTorch getters + autograd around the standard "
                   "entry vs GaussianRasterizer.forward_raw")
    return out


def run_train_iteration_compare(torch, dev, t, rs, W, H, GaussianRasterizer):
    """One whole training iteration on this workload, as train.py:90-142 strings it together, both ways:
    (reference formulation) PyTorch getters -> rasterizer (standard entry) -> torch L1 + D-SSIM -> autograd ->
    indexed densification statistics -> torch.optim.Adam over six groups;
    (fused) forward_raw -> l1_dssim_loss -> backward -> densification_stats -> FusedAdam.
    The rasterizer kernels are the same in both; the difference is N1 + N2 + N3 of SURVEY 8f."""
    import math
    from mvs_gaussian_splatting_b200.loss import l1_dssim_loss
    from mvs_gaussian_splatting_b200.optim import FusedAdam, densification_stats
    F = torch.nn.functional
    sh = t["shs"]
    raw = dict(xyz=t["means3D"], f_dc=sh[:, :1].contiguous(), f_rest=sh[:, 1:].contiguous(),
               opacity=torch.logit(t["opacities"].clamp(1e-6, 1 - 1e-6)), scaling=torch.log(t["scales"]),
               rotation=t["rotations"] * 1.7)
    lrs = dict(xyz=0.00016, f_dc=0.0025, f_rest=0.0025 / 20.0, opacity=0.05, scaling=0.005, rotation=0.001)
    gen = torch.Generator(device="cpu").manual_seed(5)
    gt = torch.rand(3, H, W, generator=gen).to(dev)
    R = GaussianRasterizer(rs)
    P = raw["xyz"].shape[0]
    lam = 0.2
    out = {}
    for name, fused in (("torch_ms", False), ("fused_ms", True)):
        ps = {k: v.detach().clone().requires_grad_(True) for k, v in raw.items()}
        opt = (FusedAdam if fused else torch.optim.Adam)([dict(params=[ps[k]], lr=lrs[k], name=k) for k in ps],
                                                         lr=0.0, eps=1e-15)
        max_radii2D = torch.zeros(P, device=dev)
        accum, denom = torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev)

        def iteration():
            m2 = torch.zeros_like(ps["xyz"], requires_grad=True)
            if fused:
                image, radii = R.forward_raw(ps["xyz"], m2, ps["f_dc"], ps["f_rest"], ps["opacity"], ps["scaling"],
                                             ps["rotation"])
                loss = l1_dssim_loss(image, gt, lam)
            else:
                image, radii = R(means3D=ps["xyz"], means2D=m2, shs=torch.cat((ps["f_dc"], ps["f_rest"]), dim=1),
                                 opacities=torch.sigmoid(ps["opacity"]), scales=torch.exp(ps["scaling"]),
                                 rotations=F.normalize(ps["rotation"]))
                g = torch.tensor([math.exp(-(i - 5) ** 2 / (2 * 1.5 ** 2)) for i in range(11)])
                g = (g / g.sum()).unsqueeze(1)
                win = g.mm(g.t()).float().expand(3, 1, 11, 11).contiguous().to(dev)
                conv = lambda a: F.conv2d(a, win, padding=5, groups=3)
                mu1, mu2 = conv(image), conv(gt)
                mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
                s1, s2, s12 = conv(image * image) - mu1_sq, conv(gt * gt) - mu2_sq, conv(image * gt) - mu12
                m = ((2 * mu12 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1_sq + mu2_sq + 1e-4) * (s1 + s2 + 9e-4))
                loss = (1.0 - lam) * torch.abs(image - gt).mean() + lam * (1.0 - m.mean())
            loss.backward()
            with torch.no_grad():
                if fused:
                    densification_stats(radii, m2.grad, max_radii2D, accum, denom)
                else:
                    vis = radii > 0
                    max_radii2D[vis] = torch.max(max_radii2D[vis], radii[vis])
                    accum[vis] += torch.norm(m2.grad[vis, :2], dim=-1, keepdim=True)
                    denom[vis] += 1
                opt.step()
                opt.zero_grad(set_to_none=True)

        for _ in range(3):
            iteration()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            iteration()
        e1.record()
        torch.cuda.synchronize()
        out[name] = e0.elapsed_time(e1) / 10
        del opt, ps
    out["note"] = ("one training iteration (render, L1 + D-SSIM, backward, densification statistics, Adam) from raw "
                   "leaf parameters: PyTorch ops around the rasterizer as the reference's train.py has them vs the "
                   "fused entries (forward_raw, l1_dssim_loss, densification_stats, FusedAdam)")
    return out


def run_ply_io(torch, dev, t):
    """the model of this workload written to and read from a Gaussian .ply (62 binary32 per row).
    Wall-clock of save_ply / load_ply (file system included; page cache warm for the read) and the
    device-side (de)interleave kernels alone (CUDA events; 484 algorithmic bytes per Gaussian)."""
    import tempfile
    import time
    from mvs_gaussian_splatting_b200 import _native, ply
    sh = t["shs"]
    args = (t["means3D"], sh[:, :1].contiguous(), sh[:, 1:].contiguous(), t["opacities"], t["scales"], t["rotations"])
    P = args[0].shape[0]
    out = {"P": P, "file_MB": round(P * 62 * 4 / 1e6, 1)}
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "point_cloud.ply")
        ply.save_ply(path, *args)                                           # warm-up (pinned allocation, page cache)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ply.save_ply(path, *args)
        out["save_ms"] = 1e3 * (time.perf_counter() - t0)
        t0 = time.perf_counter()
        got = ply.load_ply(path, dev, 3)
        torch.cuda.synchronize()
        out["load_ms"] = 1e3 * (time.perf_counter() - t0)
        out["round_trip_exact"] = bool(all(torch.equal(a.reshape(b.shape), b) for a, b in
                                           zip(args, (got["xyz"], got["features_dc"], got["features_rest"],
                                                      got["opacity"], got["scaling"], got["rotation"]))))
    cols, _ = ply.resolve_columns(ply.attribute_names(45), 3)
    rows = torch.empty(P * 62, device=dev)
    import ctypes
    ptr = lambda x: ctypes.c_void_p(x.data_ptr())
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    L = _native.lib()
    outs = [torch.empty_like(a) for a in args]
    for name, fn in (("pack_kernel_ms", lambda: L.gsr_ply_pack(P, ctypes.byref(cols), *[ptr(a) for a in args], ptr(rows), stream)),
                     ("unpack_kernel_ms", lambda: L.gsr_ply_unpack(P, ctypes.byref(cols), ptr(rows), *[ptr(a) for a in outs], stream))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        out[name] = e0.elapsed_time(e1) / 10
        out[name.replace("_ms", "_gbps")] = 484.0 * P / (out[name] * 1e-3) / 1e9
    out["note"] = ("save_ply / load_ply of this model (one read or write of the whole body through pinned memory, one "
                   "H2D/D2H copy, one (de)interleave launch); kernel lines are the launches alone")
    return out


def run_loss_compare(torch, dev, W, H):
    """the photometric loss between forward and backward, (1 - lambda) L1 + lambda (1 - SSIM)
    with lambda = 0.2 on a [3,H,W] image, value + gradient w.r.t. the rendered image — the torch formulation
    the reference uses (utils/loss_utils.py:17-63: window rebuilt and uploaded per call, five zero-padded
    grouped 11x11 conv2d, elementwise map, autograd) against l1_dssim_loss (two launches).  Algorithmic
    bytes of the fused pair: 44 B per pixel-channel."""
    import math
    from mvs_gaussian_splatting_b200.loss import l1_dssim_loss
    F = torch.nn.functional
    gen = torch.Generator(device="cpu").manual_seed(3)
    gt = torch.rand(3, H, W, generator=gen).to(dev)
    img = (gt + 0.1 * torch.randn(3, H, W, generator=gen).to(dev)).clamp(0, 1)
    lam = 0.2

    def torch_step():
        x = img.detach().requires_grad_(True)
        g = torch.tensor([math.exp(-(i - 5) ** 2 / (2 * 1.5 ** 2)) for i in range(11)])
        g = (g / g.sum()).unsqueeze(1)
        win = g.mm(g.t()).float().expand(3, 1, 11, 11).contiguous().to(dev)
        conv = lambda a: F.conv2d(a, win, padding=5, groups=3)
        mu1, mu2 = conv(x), conv(gt)
        mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
        s1, s2, s12 = conv(x * x) - mu1_sq, conv(gt * gt) - mu2_sq, conv(x * gt) - mu12
        m = ((2 * mu12 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1_sq + mu2_sq + 1e-4) * (s1 + s2 + 9e-4))
        loss = (1.0 - lam) * torch.abs(x - gt).mean() + lam * (1.0 - m.mean())
        loss.backward()
        return loss

    def fused_step():
        x = img.detach().requires_grad_(True)
        loss = l1_dssim_loss(x, gt, lam)
        loss.backward()
        return loss

    out = {}
    for name, fn in (("torch_loss_ms", torch_step), ("fused_loss_ms", fused_step)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        out[name] = e0.elapsed_time(e1) / 10
    out["fused_gbps"] = 44.0 * 3 * H * W / (out["fused_loss_ms"] * 1e-3) / 1e9
    out["note"] = ("loss value + gradient w.r.t. the rendered image at this resolution: torch ops as "
                   "utils/loss_utils.py forms them vs l1_dssim_loss (gsr_loss_forward + gsr_loss_backward)")
    return out


def run_optimizer_compare(torch, dev, t):
    """the optimizer step that follows backward, on the reference's six parameter groups
    (scene/gaussian_model.py:245-262: Adam, lr per group, eps 1e-15) at this workload's sizes —
    torch.optim.Adam as the reference constructs it against FusedAdam (one launch for all groups).  59 floats
    per Gaussian x 4 arrays (param, grad, exp_avg, exp_avg_sq): 7 x 4 B x 59 = 1652 algorithmic bytes/Gaussian."""
    from mvs_gaussian_splatting_b200.optim import FusedAdam
    sh = t["shs"]
    shapes = dict(xyz=t["means3D"], f_dc=sh[:, :1], f_rest=sh[:, 1:], opacity=t["opacities"], scaling=t["scales"],
                  rotation=t["rotations"])
    lrs = dict(xyz=0.00016, f_dc=0.0025, f_rest=0.0025 / 20.0, opacity=0.05, scaling=0.005, rotation=0.001)
    out = {}
    floats = 0
    for name, cls in (("torch_adam_ms", torch.optim.Adam), ("fused_adam_ms", FusedAdam)):
        ps = {k: v.detach().clone().contiguous().requires_grad_(True) for k, v in shapes.items()}
        for p in ps.values():
            p.grad = torch.randn_like(p) * 1e-3
        floats = sum(p.numel() for p in ps.values())
        opt = cls([dict(params=[ps[k]], lr=lrs[k], name=k) for k in ps], lr=0.0, eps=1e-15)
        for _ in range(3):
            opt.step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            opt.step()
        e1.record()
        torch.cuda.synchronize()
        out[name] = e0.elapsed_time(e1) / 10
        del opt, ps
    out["fused_gbps"] = 7 * 4 * floats / (out["fused_adam_ms"] * 1e-3) / 1e9
    out["note"] = ("Adam step over the reference's six parameter groups at this P: torch.optim.Adam (as "
                   "gaussian_model.training_setup builds it) vs FusedAdam, one gsr_adam_step launch; "
                   "fused_gbps = 28 B per float / time")
    return out


def run_e2e_host_params(args, torch, world, dev, host, rs, dL, rasterize_forward, rasterize_backward, arena):
    """Cold variant: EVERY input of the C-ABI call starts in pinned host memory (all 236 B/Gaussian)
    and every gradient is copied back to pinned host memory, each step.  PCIe-bound by construction;
    reported so the reader can see what the boundary costs when nothing is resident."""
    grad_host = torch.empty(arena.flat.numel(), dtype=torch.float32).pin_memory()

    def step():
        d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        color, radii, ctx, inp = rasterize_forward(d["means3D"], d["shs"], None, d["opacities"], d["scales"],
                                                   d["rotations"], None, rs)
        rasterize_backward(ctx, inp, rs, dL, arena=arena)
        grad_host.copy_(arena.flat, non_blocking=True)
        torch.cuda.synchronize()

    step()
    n = 5
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    dt = time.perf_counter() - t0
    h2d = sum(v.numel() * 4 for v in host.values())
    return {"value": n / dt, "unit": UNIT, "ms_per_step": 1e3 * dt / n, "steps": n, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(grad_host.numel() * 4)}


def _emit(line):
    """The ONE JSON line goes to the real stdout; everything else any library prints (NCCL's version
    banner, torchrun notices) was diverted to stderr by main()."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def run_forward_only(args):
    """Forward-only render path (render.py:32 -> render() -> rasterizer forward; quantize.py's SH-degree-0
    variant with --sh-degree 0, which passes [P,1,3] SH).  One rank; frames/s with everything resident,
    and e2e with the camera block from pinned host memory and the rendered image copied back to pinned
    host memory every frame (what render.py does before it writes the PNG)."""
    import torch
    from mvs_gaussian_splatting_b200 import GaussianRasterizationSettings, _native, last_forward_stats, rasterize_forward
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device")
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    _native.lib()
    W, H, P, D = args.width, args.height, args.points, args.sh_degree
    M = (D + 1) ** 2 if D == 0 else 16
    cam = bench_camera(0, W, H)
    g = S.make_gaussians(P, cam, sh_coeffs=M, seed=0)
    t = {k: torch.from_numpy(v).to(dev) for k, v in g.items()}
    kw = S.settings_kwargs(cam, [0.0, 0.0, 0.0], D)
    cam_host = torch.from_numpy(np.concatenate([kw["viewmatrix"].ravel(), kw["projmatrix"].ravel(),
                                                kw["campos"].ravel(), kw["bg"].ravel()]).astype(np.float32)).pin_memory()

    def settings_from(block):
        k2 = dict(kw)
        k2.update(viewmatrix=block[0:16].view(4, 4), projmatrix=block[16:32].view(4, 4), campos=block[32:35],
                  bg=block[35:38])
        return GaussianRasterizationSettings(**k2)

    rs = settings_from(cam_host.to(dev))

    def frame(settings=rs):
        return rasterize_forward(t["means3D"], t["shs"], None, t["opacities"], t["scales"], t["rotations"], None,
                                 settings)[0]

    for _ in range(max(args.warmup, 3)):
        frame()
    torch.cuda.synchronize()
    sampler = ClockSampler(dev.index or 0)
    sampler.start()
    _native.stage_timing_enable(True)
    _native.stage_timing_read()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = _native.launch_count()
    e0.record()
    for _ in range(args.steps):
        frame()
    e1.record()
    torch.cuda.synchronize()
    launches = _native.launch_count() - n0
    ms = e0.elapsed_time(e1) / args.steps
    _native.stage_timing_enable(False)
    stage_ms = {k: v[0] / args.steps for k, v in _native.stage_timing_read().items() if v[1]}
    clocks = sampler.result()
    st = last_forward_stats(dev)
    # e2e: camera H2D + image D2H every frame, double-buffered pinned image
    img_host = [torch.empty((3, H, W), dtype=torch.float32).pin_memory() for _ in range(2)]
    done = [torch.cuda.Event(), torch.cuda.Event()]

    def e2e_frame(k):
        block = cam_host.to(dev, non_blocking=True)
        color = frame(settings_from(block))
        done[k & 1].synchronize()                      # the copy that last used this pinned buffer is complete
        img_host[k & 1].copy_(color, non_blocking=True)
        done[k & 1].record()

    for k in range(3):
        e2e_frame(k)
    torch.cuda.synchronize()
    n = max(5, min(args.steps, 20))
    t0 = time.perf_counter()
    for k in range(n):
        e2e_frame(k)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    peaks = load_peaks()
    line = {"metric": "rasterizer_fwd_frames_per_s", "value": 1e3 / ms, "unit": "frames/s", "n_gpus": 1,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{P} Gaussians SH{D} (M={M}) {W}x{H} forward only", "baseline_config":
                       "BASELINE.json configs[4]", "l2": "inputs larger than L2, no flush",
                       "counts": {"P": P, "P_visible": st["num_visible"], "R": st["num_rendered"]}},
            "gpu_launches": int(launches), "clocks": clocks,
            "stages_ms": stage_ms, "stage_ms_sum": sum(stage_ms.values()),
            "roofline": {"bound": "hbm", "achieved": (52.0 * P + (67.0 + 12.0 * M) * st["num_visible"]) / (stage_ms.get("preprocess", ms) * 1e-3) / 1e9,
                         "peak": peaks["hbm_gbs"], "unit": "GB/s", "traffic": None, "stage": "preprocess",
                         "note": "per-stage times in stages_ms; blend_forward is FP32-issue bound (DESIGN.md 4)"},
            "e2e": {"value": n / dt, "unit": "frames/s", "ms_per_step": 1e3 * dt / n, "steps": n,
                    "h2d_bytes_per_step": int(cam_host.numel() * 4), "d2h_bytes_per_step": int(3 * H * W * 4),
                    "api": "rasterize_forward (GaussianRasterizer forward under no_grad), camera from pinned host "
                           "memory, image copied to pinned host memory every frame"},
            "cpu_baseline": None}
    line["roofline"]["frac"] = line["roofline"]["achieved"] / peaks["hbm_gbs"]
    _emit(line)


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)  # C-level writes to fd 1 (NCCL_DEBUG banners) must not pollute the JSON line
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.forward_only:
        run_forward_only(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
