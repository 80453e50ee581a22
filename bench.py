"""bench.py — fwd+bwd of the differentiable Gaussian-splatting rasterizer on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --steps K --warmup W    # the CPU arm (oracle port, host cores)

Workloads.  N = 1 (the headline): BASELINE.json configs[2], the one the metric is quoted on — 3 M synthetic
Gaussians, SH degree 3, 1920x1080, one view per step, forward + backward with a fixed seeded dL/dpixel.
N > 1: BASELINE.json configs[3] — 6 M Gaussians, an 8-VIEW batch per step split over the N GPUs (rank r
renders views r, r+N, ...; cameras on SURVEY 8d's radius-4 ring), Gaussian-sharded: each rank owns 1/N of
the parameters and of their gradients; the exchange is fused into the per-Gaussian kernels over NVLink peer
memory.  The same batch on ONE GPU is measured inside the N = 1 line (key `multi_view_batch`), which is the
single-GPU point of that scaling curve.  `--points / --views` select other sizes.
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
    ap.add_argument("--points", type=int, default=None,
                    help="Gaussians (default: 3 000 000 at one GPU = configs[2]; 6 000 000 at N > 1 = configs[3])")
    ap.add_argument("--views", type=int, default=None,
                    help="views per step, a multiple of the GPU count (default: 1 at one GPU; 8 at N > 1)")
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


def workload_sizes(args, world: int):
    """(P, V): defaults follow BASELINE.json — configs[2] on one GPU, configs[3] (6 M, 8-view batch) on N > 1."""
    P = args.points if args.points is not None else (3_000_000 if world == 1 else 6_000_000)
    V = args.views if args.views is not None else (1 if world == 1 else 8)
    if V < 1 or V % world:
        raise SystemExit(f"bench.py: --views {V} must be a positive multiple of --gpus {world}")
    return P, V


def workload_text(P, D, W, H, V) -> str:
    """The ONE description of the workload both arms print (the driver compares the strings)."""
    return (f"{P} Gaussians SH{D} {W}x{H} fwd+bwd, " +
            ("1 view per step (front camera)" if V == 1 else f"{V}-view batch per step (cameras on a radius-4 ring)"))


def baseline_config(P, V, W, H) -> str:
    if (P, V, W, H) == (3_000_000, 1, 1920, 1080):
        return "BASELINE.json configs[2]"
    if (P, V, W, H) == (6_000_000, 8, 1920, 1080):
        return "BASELINE.json configs[3]"
    return "custom (--points/--views/--width/--height)"


def measured_simt_peaks():
    """FP32 / MUFU peaks measured on this pool's B200 with tools/fp32_peak.cu (profiles/r02a_fp32_peak.json);
    MEASURED_PEAKS.json carries HBM and tensor-core figures only."""
    p = os.path.join(ROOT, "profiles", "r02a_fp32_peak.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return dict(ffma_tflops=float(d["ffma_reg_tflops"]["best"]), ffma_3src_tflops=float(d["ffma_3src_tflops"]["best"]),
                    mufu_tops=float(d["mufu_ex2_tops"]["best"]), source="measured: tools/fp32_peak.cu on B200 "
                    "(profiles/r02a_fp32_peak.json; FFMA with two distinct source registers; with three distinct "
                    "sources the register file sustains " + f"{float(d['ffma_3src_tflops']['best']):.1f} TFLOP/s)")
    except (OSError, KeyError, ValueError):
        nominal = N_SM * FP32_LANES_PER_SM * 2 * 1965e6 / 1e12
        return dict(ffma_tflops=nominal, ffma_3src_tflops=None, mufu_tops=None,
                    source="nominal 148 SM x 128 lanes x 2 x 1.965 GHz (no measured SIMT figure found)")


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

    def forward_only(self, budget_s=12.0):
        """Forward-only frame on the host cores (configs[4]): preprocess + binning in full, the blend on every
        s-th tile row when the budget requires it (time scaled by rows)."""
        O, g, cam, th = self.O, self.g, self.cam, self.threads
        W, H = cam.width, cam.height
        bg = np.zeros(3, np.float32)
        t0 = time.perf_counter()
        pre = O.preprocess(g["means3D"], g["scales"], g["rotations"], g["opacities"], g["shs"], cam.world_view_transform,
                           cam.full_proj_transform, cam.camera_center, W, H, cam.tanfovx, cam.tanfovy, self.D, threads=th)
        binned = O.bin_instances(pre, W, H)
        t1 = time.perf_counter()
        # calibrate the blend on 1/16 of the tile rows, then size the sample for the budget
        tr = self.rows(16)
        O.blend_forward(pre, binned, bg, W, H, threads=th, tile_ranges=tr)
        t_cal = (time.perf_counter() - t1) * self.gy / len(tr)
        stride = 1
        while stride < self.gy and t_cal * len(self.rows(stride)) / self.gy > budget_s:
            stride += 1
        tr = None if stride == 1 else self.rows(stride)
        t2 = time.perf_counter()
        O.blend_forward(pre, binned, bg, W, H, threads=th, tile_ranges=tr)
        t3 = time.perf_counter()
        frac = 1.0 if stride == 1 else len(tr) / self.gy
        t_frame = (t1 - t0) + (t3 - t2) / frac
        return {"value": 1.0 / t_frame, "unit": "frames/s", "ms_per_frame": 1e3 * t_frame, "cores": th,
                "kind": "port" if stride == 1 else "port-extrapolated",
                "sample": "preprocess + binning of all Gaussians; blend forward on " +
                          ("all tile rows" if stride == 1 else f"every {stride}th tile row ({len(tr)}/{self.gy}), time scaled by rows")}

    def describe(self, stride):
        if stride == 1:
            return "full view: all Gaussians, all tiles"
        return (f"all Gaussians through preprocess/binning/preprocess-backward; blend fwd+bwd on every "
                f"{stride}th tile row ({len(self.rows(stride))}/{self.gy} rows), blend time scaled by rows")


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # under torchrun only rank 0 runs the CPU arm
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    W, H, D = args.width, args.height, args.sh_degree
    P, V = workload_sizes(args, max(world, args.gpus))
    cam = S.batch_cameras(V, W, H)[0]            # one view of the batch: views/s = 1 / seconds per view
    g = S.make_gaussians(P, S.front_camera(W, H), seed=0)
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
    kind = "port" if stride == 1 else "port-extrapolated"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_text(P, D, W, H, V), "baseline_config": baseline_config(P, V, W, H),
                   "arm": "CPU oracle port (oracle/gs_oracle.c), all host cores; the reference's own CUDA rasterizer is "
                          "an un-vendored submodule and is not installed on the box (tools/probe_reference_cuda.py); "
                          "one view of the batch is timed, views/s = 1 / seconds per view"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": th, "kind": kind, "sample": arm.describe(stride)},
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
    keybits = 32 + max(1, math.ceil(math.log2(max(tiles / max(counts.get("views_rendered_by_this_rank", 1), 1), 2))))
    tile_key_bytes = 2.0 if keybits <= 48 else 4.0
    hbm = {
        "preprocess": 52.0 * P + 259.0 * Pv,
        # pass 1 reads P keys and writes the P_visible (key, index) pairs it keeps, passes 2-4 move P_visible pairs
        # (8 B read + 8 B write); the counts + histograms come from the persistent preprocess kernel on the single-view
        # path, from one more read of depths and tiles_touched (8 P) on the multi-view / sharded path
        "depth_sort": (4.0 if counts.get("depth_counts_in_preprocess") else 12.0) * P + 56.0 * Pv,
        # prefix sum + emission in one kernel: order, radii, means2D read per visible Gaussian; one
        # (tile id, index) entry written per instance (2 + 4 B up to 65536 tiles)
        "emit_keys": 16.0 * Pv + (tile_key_bytes + 4.0) * R,
        "tile_ranges": 12.0 * tiles,     # per-tile counts read, [start, end) written
        # tile-bit sort only (the depth order comes from depth_sort): passes x (entry read + entry written)
        "radix_sort": 2.0 * (tile_key_bytes + 4.0) * math.ceil((keybits - 32) / 8) * R,
        "preprocess_backward": 303.0 * Pv + 240.0 * P,
    }
    flops = {
        "blend_forward": FLOP_PER_EVAL * E + FLOP_PER_BLEND_FWD * C,
        "blend_backward": FLOP_PER_EVAL * En + FLOP_PER_BLEND_BWD * C,
    }
    simt = measured_simt_peaks()
    fp32_peak = simt["ffma_tflops"]   # TFLOP/s, FMA = 2 flops
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
                         "peak_source": simt["source"],
                         "note": "achieved counts the REFERENCE algorithm's flops per launch (18 per evaluated pair + "
                                 "8 / 73 per blended pair); culling never issues most of them, so this is a useful-work "
                                 "rate — pipe utilisation is the ncu issue-slot figure in `issue_slot_pct`",
                         "issue_slot_pct": ncu_issue.get(name)}
    return out


def load_ncu_issue():
    """Issue-slot utilisation (smsp__issue_active, % of peak) of the two blend kernels from the committed
    ncu --set full capture (profiles/*_ncu_full_summary.md); {} when none is recorded."""
    p = os.path.join(ROOT, "profiles", "issue_slots.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


ncu_issue = load_ncu_issue() if os.path.isdir(os.path.join(ROOT, "profiles")) else {}


def load_traffic():
    """dram bytes per launch from the committed ncu --set full capture, if one is recorded."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return {}


def clock_hz(torch, dev):
    """SM clock of `dev` in Hz (torch.cuda._sleep counts cycles)."""
    try:
        return 1e3 * float(torch.cuda.get_device_properties(dev).clock_rate)
    except Exception:
        return 1.9e9


def _event_stats(events, world, dist, torch, dev):
    """Per-step device times from K+1 events: (mean, median, p90) in ms, each the MAX over ranks."""
    per = np.array([events[i].elapsed_time(events[i + 1]) for i in range(len(events) - 1)], np.float64)
    vals = [float(events[0].elapsed_time(events[-1])) / max(len(per), 1), float(np.median(per)),
            float(np.quantile(per, 0.9))]
    if world > 1:
        tt = torch.tensor(vals, device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        vals = [float(x) for x in tt.tolist()]
    return vals + [[round(float(x), 3) for x in per]]


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

    W, H, D = args.width, args.height, args.sh_degree
    P, V = workload_sizes(args, world)
    M = 16
    warmup = max(args.warmup, 3 if world == 1 else 20)   # the double-buffered peer exchange needs a longer run-in
    g = S.make_gaussians(P, S.front_camera(W, H), sh_coeffs=M, seed=0)   # same scene on every rank
    cam_list = S.batch_cameras(V, W, H)
    host = {k: torch.from_numpy(v).pin_memory() for k, v in g.items()}
    t = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
    multi = V > 1 or world > 1            # a batch of views: the Gaussian-sharded step (world 1: "local" transport)
    if multi and not args.multi.startswith("sharded") and V != world:
        raise SystemExit("bench.py: --multi allreduce renders one view per rank (use --views equal to --gpus)")

    def cam_block(c):
        kw = S.settings_kwargs(c, [0.0, 0.0, 0.0], D)
        return np.concatenate([kw["viewmatrix"].ravel(), kw["projmatrix"].ravel(), kw["campos"].ravel(),
                               kw["bg"].ravel()]).astype(np.float32)

    kw0 = S.settings_kwargs(cam_list[0], [0.0, 0.0, 0.0], D)

    def settings_from(block, v=0):
        k2 = S.settings_kwargs(cam_list[v], [0.0, 0.0, 0.0], D) if v else dict(kw0)
        k2.update(viewmatrix=block[0:16].view(4, 4), projmatrix=block[16:32].view(4, 4), campos=block[32:35],
                  bg=block[35:38])
        return GaussianRasterizationSettings(**k2)

    cam_host = torch.from_numpy(np.concatenate([cam_block(c) for c in cam_list])).pin_memory()   # 38 floats per view
    cam_dev = cam_host.to(dev)
    cams_all = [settings_from(cam_dev[38 * v:38 * v + 38], v) for v in range(V)]
    rs = cams_all[rank % V]
    rng = np.random.Generator(np.random.PCG64(1234))
    dL_np = rng.normal(size=(3, H, W)).astype(np.float32)
    dL = torch.from_numpy(dL_np).to(dev)
    sharded = multi and args.multi.startswith("sharded")
    arena = None if sharded else GradientArena(P, M, dev)
    sr = slices = dL_local = None
    if sharded:
        from mvs_gaussian_splatting_b200.sharded import ShardedViewRasterizer
        sr = ShardedViewRasterizer(P, M, W, H, dev, exchange="nccl" if args.multi == "sharded-nccl" else "peer",
                                   n_views=V)
        rows = sr.plan.rows(rank)
        slices = {k: t[k][rows].contiguous() for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        dL_local = dL.unsqueeze(0).expand(sr.L, 3, H, W).contiguous()

    def step():
        if sharded:
            sr.forward(slices, cams_all)
            sr.backward(dL_local)
            return sr._state[0][0]
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

    for _ in range(warmup):
        ctx = step()
    # NVML is initialised BEFORE the barrier: nvmlInit takes milliseconds that differ from rank to rank, and host skew
    # after the barrier lands in the first timed step of every rank (the peers wait for the slowest at the exchange)
    sampler = ClockSampler(local)
    events = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    align_token = torch.zeros(1, device=dev)
    if world > 1:
        dist.all_reduce(align_token)   # communicator warm for the rendezvous below
    barrier()
    # ---- headline: K steps, stage timing OFF, one event per step boundary on the launching stream ----
    sampler.start()
    n0 = _native.launch_count()
    # The barrier above leaves every launch queue empty and the ranks' hosts skewed: the first timed step would then
    # measure Python's launch latency (and, with peers, the slowest host) instead of the device (measured: +0.4 ms at
    # 1 GPU, +2.6 ms at 8).  A device-side delay OUTSIDE the timed region (before the first event) lets the hosts run
    # one step ahead, as they do in every later step; stated in config.timing.
    prime_ms = 2.0 if world == 1 else 8.0
    torch.cuda._sleep(int(prime_ms * 1e-3 * (clock_hz(torch, dev))))
    if world > 1:
        # ... and a device-side rendezvous behind the delay: a host that was descheduled for a millisecond after the
        # barrier starts its delay late, and without this its peers' first step would wait for it at the exchange
        dist.all_reduce(align_token)
    events[0].record()
    for i in range(args.steps):
        ctx = step()
        events[i + 1].record()
    barrier()
    launches = _native.launch_count() - n0
    clocks = sampler.result()
    ms_step, ms_median, ms_p90, ms_all = _event_stats(events, world, dist, torch, dev)
    value = V * 1e3 / ms_step
    # ---- second pass: the same K steps with the library's per-stage event pairs switched on ----
    _native.stage_timing_enable(True)
    _native.stage_timing_read()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        ctx = step()
    e1.record()
    barrier()
    ms_stage_pass = e0.elapsed_time(e1) / args.steps
    _native.stage_timing_enable(False)
    stages = _native.stage_timing_read()

    # ---- counts for the algorithmic cost model (outside the timed region) ----
    gx, gy = (W + 15) // 16, (H + 15) // 16
    L_views = sr.L if sharded else 1
    if sharded:   # sums over this rank's views of the step
        cts = sr._state[0]
        pc = [blend_pair_counts(c) for c in cts]
        ev, bl, ev_ours, warp_iters = (sum(x[i] for x in pc) for i in range(4))
        nvis = sum(int((c.radii > 0).sum().item()) for c in cts)
        nren = sum(int(c.num_rendered) for c in cts)
        ncontrib = sum(int(c.array("n_contrib").long().sum().item()) for c in cts)
        per_view_R = [int(c.num_rendered) for c in cts]
    else:
        st = last_forward_stats(dev)
        ev, bl, ev_ours, warp_iters = blend_pair_counts(ctx)
        nvis, nren = st["num_visible"], st["num_rendered"]
        ncontrib = int(ctx.array("n_contrib").long().sum().item())
        per_view_R = [int(nren)]
    # per RANK and step: this rank preprocesses its slice under all V views and renders L views of all P Gaussians
    counts = dict(P=P * L_views, P_visible=nvis, R=nren, tiles=gx * gy * L_views, pairs_evaluated=ev,
                  pairs_blended=bl, pairs_evaluated_after_culling=ev_ours, warp_splat_iterations=warp_iters,
                  sum_n_contrib=ncontrib, views_rendered_by_this_rank=L_views, instances_per_view=per_view_R,
                  depth_counts_in_preprocess=not sharded)
    stage_ms = {k: v[0] / max(args.steps, 1) for k, v in stages.items() if v[1]}   # per step (a stage may launch N times)
    peaks = load_peaks()
    roofs = stage_rooflines(stage_ms, counts, peaks, clocks)
    traffic = load_traffic() if (P, V, W, H) == (3_000_000, 1, 1920, 1080) else {}
    for k, r in roofs.items():
        r["traffic"] = traffic.get(k)
    dominant = max(stage_ms, key=stage_ms.get)
    per_rank = None
    if world > 1:   # which rank sets the pace: the ranks render different views (different R, different blend cost)
        mine = dict(rank=rank, views=(sr.local_views() if sharded else [rank]), num_rendered=per_view_R,
                    stage_ms_sum=round(sum(stage_ms.values()), 4),
                    blend_ms=round(stage_ms.get("blend_forward", 0.0) + stage_ms.get("blend_backward", 0.0), 4))
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        per_rank = gathered

    # ---- e2e: the public API with host-side per-step inputs ----
    e2e = run_e2e(args, torch, dist, world, dev, t, cam_host, settings_from, dL_np, GaussianRasterizer, V,
                  sharded=(sr, slices, rank) if sharded else None)
    e2e_hp = run_e2e_host_params(args, torch, world, dev, host, rs, dL, rasterize_forward, rasterize_backward,
                                 arena) if (world == 1 and not sharded) else None

    exchange = None
    if sharded:
        remote = (world - 1) / world
        exchange = {"transport": sr.exchange, "forward_bytes_per_gaussian_view": 52, "backward_bytes_per_gaussian_view": 48,
                    "nvlink_bytes_per_rank_step": int((52 + 48) * (P / world) * V * remote),
                    "note": "rows written straight into / accumulators read straight from the rendering rank's workspace by the "
                            "per-Gaussian kernels (peer memory); each rank owns 1/N of the parameters and gradients"}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": warmup, "ms_per_step": ms_step, "ms_per_step_median": ms_median, "ms_per_step_p90": ms_p90,
        "ms_per_step_rank0": ms_all,
        "ms_per_view": ms_step / V, "higher_is_better": True,
        # the sharded step splits ONE fixed V-view batch over the N ranks: strong scaling at every N > 1 (the N = 1
        # headline is configs[2], a different workload; config.n1_same_workload_views_per_s is this batch on one GPU)
        "scaling": "strong" if (sharded and world > 1) else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_text(P, D, W, H, V), "baseline_config": baseline_config(P, V, W, H),
                   "l2": f"inputs ({236 * P / 1e6:.0f} MB) larger than L2, no flush",
                   "parallelism": (f"{V} views over {world} GPU(s), Gaussian-sharded" if sharded else
                                   f"views x{world}" + (" + NCCL all-reduce of 59 f32/Gaussian gradients" if world > 1 else "")),
                   "exchange": exchange, "counts": counts,
                   "timing": "value: K steps with per-stage timing OFF, CUDA events per step boundary; after the barrier + synchronize "
                             "and before the first event the stream holds a device-side delay (2 ms; 8 ms with peers) so that "
                             "the launch queue is primed as in every later step instead of timing Python's launch latency "
                             "on an empty queue (with peers, a one-word all-reduce behind the delay aligns the ranks' first event); stages: a second pass of K steps with the "
                             "library's event pairs on (stage_pass_ms_per_step)"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofs.get(dominant),
        "dominant_stage": dominant,
        "stages": roofs,
        "stage_ms_sum": sum(stage_ms.values()),
        "stage_pass_ms_per_step": ms_stage_pass,
        "e2e": e2e,
    }
    if e2e_hp is not None:
        line["e2e_host_params"] = e2e_hp
    if per_rank is not None:
        line["per_rank"] = per_rank
    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        try:
            import probe_reference_cuda
            line["reference_cuda"] = probe_reference_cuda.summary()
        except Exception as e:   # the probe must never break the bench line
            line["reference_cuda"] = f"probe failed: {e}"
    if sharded and world > 1 and not args.no_extras:
        # (1) parity of THIS run's workload on THIS box: every rank renders the batch with the single-GPU entry points
        # on the full parameter set and with the sharded step on its slice (tools/sharded_check.run_check: images,
        # radii, sorted lists bit-exact; slice gradients / statistics to 5e-5 relative L1).  A mismatch fails the run.
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import sharded_check
        del sr, slices, dL_local
        torch.cuda.empty_cache()
        ok, worst = sharded_check.run_check(P, W, H, V, dev, rank, world, "nccl" if args.multi == "sharded-nccl" else "peer")
        flag = torch.tensor([0 if ok else 1], device=dev)
        dist.all_reduce(flag)
        parity_ok = flag.item() == 0
        line["config"]["sharded_parity"] = (f"ok (all {world} ranks; worst slice-gradient rel_l1 on rank 0 {worst:.1e})"
                                            if parity_ok else f"FAIL on {int(flag.item())} rank(s)")
        # (2) the same batch on ONE GPU of this box (rank 0, world-1 "local" transport; the other ranks wait): the
        # like-for-like single-GPU point of the scaling curve, as scalars the driver's record keeps
        torch.cuda.empty_cache()
        if rank == 0:
            one = run_multi_view_batch(torch, dev, D, W, H, 8, P=P, V=V, t=t, standalone=True)
            line["config"]["n1_same_workload_views_per_s"] = one["value"]
            line["config"]["n1_same_workload_ms_per_step"] = one["ms_per_step"]
            line["config"]["efficiency_vs_n1_same_workload"] = value / (world * one["value"])
            if e2e and e2e.get("value"):
                line["config"]["e2e_efficiency_vs_n1_same_workload_device_timed"] = e2e["value"] / (world * one["value"])
        dist.barrier()
        if not parity_ok:
            if rank == 0:
                _emit(line)
            dist.destroy_process_group()
            raise SystemExit("bench.py: the sharded step does not match the single-GPU path on this workload")
    if world == 1 and not multi and not args.no_extras:
        line["multi_view_batch"] = run_multi_view_batch(torch, dev, D, W, H, min(args.steps, 5))
        line["config"]["n1_multi_view_batch_views_per_s"] = line["multi_view_batch"]["value"]
        line["forward_only"] = run_forward_only_extra(torch, dev, min(args.steps, 10), not args.no_cpu_baseline)
        for name in ("sh3", "sh0"):    # configs[4] as scalars the driver's record keeps
            line["config"][f"forward_only_fps_{name}"] = line["forward_only"][name]["value"]
        line["raw_entry"] = run_raw_compare(args, torch, dev, t, rs, dL, GaussianRasterizer)
        line["optimizer_step"] = run_optimizer_compare(torch, dev, t)
        line["loss_step"] = run_loss_compare(torch, dev, W, H)
        line["ply_io"] = run_ply_io(torch, dev, t)
        line["train_iteration"] = run_train_iteration_compare(torch, dev, t, rs, W, H, GaussianRasterizer)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(g, cam_list[0], D, dL_np)
        line["cpu_baseline_2d"] = cpu_baseline_2d()
    if rank == 0:
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_multi_view_batch(torch, dev, D, W, H, steps, P=6_000_000, V=8, t=None, standalone=False):
    """BASELINE.json configs[3] on ONE GPU — the single-GPU point of the multi-GPU scaling curve: the same
    Gaussian-sharded step with world 1 ("local" transport): the multi-view per-Gaussian kernels read the 6 M
    parameters once for the 8 views, every view is binned / blended in turn, gradients are summed over the
    batch inside preprocess_backward.  Device-timed, K steps after 3 warm-up steps."""
    from mvs_gaussian_splatting_b200 import GaussianRasterizationSettings
    from mvs_gaussian_splatting_b200.sharded import ShardedViewRasterizer
    M = 16
    own_t = t is None
    if own_t:
        g = S.make_gaussians(P, S.front_camera(W, H), sh_coeffs=M, seed=0)
        t = {k: torch.from_numpy(v).to(dev) for k, v in g.items()}
    cams = []
    for c in S.batch_cameras(V, W, H):
        kw = S.settings_kwargs(c, [0.0, 0.0, 0.0], D)
        for k in ("bg", "viewmatrix", "projmatrix", "campos"):
            kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k])).to(dev)
        cams.append(GaussianRasterizationSettings(**kw))
    sr = ShardedViewRasterizer(P, M, W, H, dev, n_views=V, standalone=standalone)
    gen = torch.Generator(device="cpu").manual_seed(1234)
    dL = torch.randn(3, H, W, generator=gen).to(dev).unsqueeze(0).expand(V, 3, H, W).contiguous()
    for _ in range(3):
        sr.forward(t, cams)
        sr.backward(dL)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        sr.forward(t, cams)
        sr.backward(dL)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    out = {"value": V * 1e3 / ms, "unit": UNIT, "ms_per_step": ms, "ms_per_view": ms / V, "steps": steps, "n_gpus": 1,
           "workload": workload_text(P, D, W, H, V), "baseline_config": baseline_config(P, V, W, H),
           "instances_per_view": [int(c.num_rendered) for c in sr._state[0]],
           "note": "the N = 1 point of the configs[3] scaling curve: bench.py --gpus N (N > 1) runs this same batch "
                   "split over N GPUs; efficiency(N) = value(N) / (N x this value)"}
    del sr
    if own_t:
        del t
    torch.cuda.empty_cache()
    return out


def run_forward_only_extra(torch, dev, steps, with_cpu):
    """BASELINE.json configs[4] inside the default line (so the driver records it): forward-only render of 6 M
    Gaussians at 3840x2160, SH degree 3 and the SH-degree-0 variant, frames/s device-timed; plus the CPU oracle's
    forward-only time on the host cores (bounded sample of tile rows)."""
    from mvs_gaussian_splatting_b200 import GaussianRasterizationSettings, last_forward_stats, rasterize_forward
    P, W, H = 6_000_000, 3840, 2160
    cam = S.front_camera(W, H)
    g = S.make_gaussians(P, cam, sh_coeffs=16, seed=0)
    out = {"workload": f"{P} Gaussians {W}x{H} forward only (render.py path)", "baseline_config": "BASELINE.json configs[4]"}
    t = {k: torch.from_numpy(v).to(dev) for k, v in g.items()}
    for name, Dv in (("sh3", 3), ("sh0", 0)):
        kw = S.settings_kwargs(cam, [0.0, 0.0, 0.0], Dv)
        for k in ("bg", "viewmatrix", "projmatrix", "campos"):
            kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k])).to(dev)
        rs = GaussianRasterizationSettings(**kw)
        shs = t["shs"] if Dv else t["shs"][:, :1].contiguous()     # quantize.py's variant passes [P,1,3]
        frame = lambda: rasterize_forward(t["means3D"], shs, None, t["opacities"], t["scales"], t["rotations"], None, rs)
        for _ in range(3):
            frame()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            frame()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        st = last_forward_stats(dev)
        out[name] = {"value": 1e3 / ms, "unit": "frames/s", "ms_per_frame": ms, "steps": steps,
                     "P_visible": st["num_visible"], "R": st["num_rendered"], "sh_coeffs": int(shs.shape[1])}
    del t
    torch.cuda.empty_cache()
    if with_cpu:
        th = host_threads()
        arm = CpuArm(g, cam, 3, None, th)
        out["cpu_baseline"] = arm.forward_only(budget_s=12.0)
    return out


def cpu_baseline_2d():
    """north_star's CPU baseline: the reference's pure-PyTorch dense 2-D splat compositing
    (2D-Gaussian-Splatting-main/2d_gaussian_splatting.py:44-123, restated in oracle/dense_splat_2d.py) timed on the
    host cores in this same run: its own configuration (config.yml: 1000 Gaussians, 256x256, kernel 101) and a
    scaled slice at 1920x1080.  A different function from the 3-D rasterizer (additive, unsorted): a throughput
    baseline in ns per splat*pixel, nothing more."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import dense_splat_2d
    th = host_threads()
    a = dense_splat_2d.time_dense_splat_2d(1000, 256, 256, reps=1, threads=th)
    b = dense_splat_2d.time_dense_splat_2d(16, 1080, 1920, reps=1, threads=th)
    # what a dense pass over configs[2]'s visible splats would cost at that rate (it is O(N x pixels))
    est = b["fwd_bwd_ns_per_splat_pixel"] * 1e-9 * 2_285_341 * 1920 * 1080
    return {"kind": "reference-restated", "cores": th, "config1": a, "slice_1080p": b,
            "extrapolated_fwd_bwd_s_per_view_config2": est,
            "note": "dense compositing touches every pixel for every splat; the extrapolation multiplies the measured "
                    "1080p rate by configs[2]'s 2.29 M visible splats (labelled extrapolated, SURVEY 8d)"}


def cpu_baseline(g, cam, D, dL_np, budget_s=30.0):
    th = host_threads()
    arm = CpuArm(g, cam, D, dL_np, th)
    cal = arm.step(1)                      # warm-up + calibration (untimed)
    stride = 1
    while stride < arm.gy and 2 * arm_estimate(cal, stride, arm.gy) > budget_s:
        stride += 1
    ts = [arm.step(stride)["t_view_s"] for _ in range(2)]
    t = float(np.mean(ts))
    return {"value": 1.0 / t, "unit": UNIT, "cores": th, "kind": "port" if stride == 1 else "port-extrapolated",
            "sample": arm.describe(stride), "ms_per_view": 1e3 * t}


def run_e2e(args, torch, dist, world, dev, t, cam_host, settings_from, dL_np, GaussianRasterizer, V, sharded=None):
    """The call a user of the reference makes (gaussian_renderer/__init__.py:257-265 + train.py:95-113):
    per step the camera blocks of the batch and the ground-truth images of the views this rank renders come from
    PINNED HOST memory, the images are rendered through the public API (GaussianRasterizer under autograd at one
    view per step; ShardedViewRasterizer for a batch), an L1 loss is formed per view, backward runs, and the loss
    is read back to the host.  The Gaussian parameters are model state and stay in HBM, as in the reference's
    training loop."""
    from mvs_gaussian_splatting_b200.loss import l1_loss
    H, W = args.height, args.width
    params = sr = slices = None
    L = 1
    if sharded is not None:
        sr, slices, my_rank = sharded
        L = sr.L
    else:
        params = {k: t[k].detach().clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    # Ground truth travels as the dataset stores it — 8 bits per channel (the reference's readers decode PNG / JPEG
    # and divide by 255, utils/general_utils.py PILtoTorch) — and becomes float32 on the device: a quarter of the
    # PCIe bytes of a float upload (the reference itself keeps its float images resident, scene/cameras.py:39).
    gt1 = np.rint((0.5 + 0.25 * dL_np).clip(0, 1) * 255.0).astype(np.uint8)
    gt_host = torch.from_numpy(np.ascontiguousarray(np.broadcast_to(gt1, (L, 3, H, W))).copy()).pin_memory()

    # Per-step host inputs are prefetched one step ahead on a copy stream (double-buffered), as a data
    # loader does; the loss is copied to pinned memory asynchronously and read one step late, so the
    # host never stalls the launch queue.  Every copy still happens inside the timed region, once per step.
    copy_stream = torch.cuda.Stream(device=dev)
    main = torch.cuda.current_stream(dev)
    slots = [dict(block=torch.empty_like(cam_host, device=dev), gt_u8=torch.empty_like(gt_host, device=dev),
                  gt=torch.empty(gt_host.shape, dtype=torch.float32, device=dev),
                  ready=torch.cuda.Event(), free=torch.cuda.Event(), loss=torch.zeros(L).pin_memory(),
                  loss_ready=torch.cuda.Event(), h2d=(torch.cuda.Event(enable_timing=True),
                                                      torch.cuda.Event(enable_timing=True))) for _ in range(2)]
    for sl in slots:
        sl["free"].record(main)
        # the settings NamedTuples are views of the slot's device block: built once per slot, refreshed by the copy
        sl["cams"] = [settings_from(sl["block"][38 * v:38 * v + 38], v) for v in range(V)]
    state = {"k": 0, "pending": None}
    skip = set(filter(None, os.environ.get("GSR_E2E_SKIP", "").split(",")))   # diagnostic: drop pieces of the loop

    def prefetch(sl):
        if "h2d" in skip:
            return
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(sl["free"])          # the step that last used this slot is done with it
            sl["h2d"][0].record(copy_stream)
            sl["block"].copy_(cam_host, non_blocking=True)
            sl["gt_u8"].copy_(gt_host, non_blocking=True)
            sl["h2d"][1].record(copy_stream)
            sl["ready"].record(copy_stream)

    prefetch(slots[0])

    def step():
        sl = slots[state["k"] & 1]
        prefetch(slots[(state["k"] + 1) & 1])           # next step's inputs travel while this step computes
        if "h2d" not in skip:
            main.wait_event(sl["ready"])
        if "mul" not in skip:
            torch.mul(sl["gt_u8"], 1.0 / 255.0, out=sl["gt"])    # 8-bit image -> float32 on the device
        if sr is not None:
            colors, radii = sr.forward(slices, sl["cams"])
            colors.requires_grad_(True)
            losses = [l1_loss(colors[l], sl["gt"][l]) for l in range(L)]   # library loss (loss.py): one launch each way
            loss = torch.stack(losses)
            loss.sum().backward()
            sr.backward(colors.grad)
        else:
            rs = sl["cams"][int(os.environ.get("RANK", "0")) % V]
            means2D = torch.zeros_like(params["means3D"], requires_grad=True)
            color, radii = GaussianRasterizer(rs)(means3D=params["means3D"], means2D=means2D, shs=params["shs"],
                                                  opacities=params["opacities"], scales=params["scales"],
                                                  rotations=params["rotations"])
            loss = l1_loss(color, sl["gt"][0])
            loss.backward()
            if world > 1:
                for p in params.values():
                    dist.all_reduce(p.grad)
        if "d2h" not in skip:
            sl["loss"].copy_(loss.detach().reshape(L), non_blocking=True)   # D2H read of the step's result
        sl["loss_ready"].record(main)
        sl["free"].record(main)
        val = None
        if state["pending"] is not None and "d2h" not in skip:   # previous step's loss, now on the host
            state["pending"]["loss_ready"].synchronize()
            val = float(state["pending"]["loss"][0])
        state["pending"] = sl
        state["k"] += 1
        if params is not None:
            for p in params.values():
                p.grad = None                               # train.py:142 zero_grad(set_to_none=True)
        return val

    import gc
    for _ in range(10 if world == 1 else 20):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    # wall-clock: enough steps (>= 0.15 s) that one host hiccup or the empty launch queue of the first step does not
    # move the figure; the collector is paused as a training loop's inner iteration would not trigger it either
    n = max(100 if V == 1 else 30, args.steps)
    gc.collect()
    gc.disable()
    prof = None
    if os.environ.get("GSR_E2E_PROFILE") and int(os.environ.get("RANK", "0")) == 0:
        import cProfile                                   # where the host time of an e2e step goes (stderr; perturbs the figure)
        prof = cProfile.Profile()
        prof.enable()
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gc.enable()
    h2d_ms = 0.0 if "h2d" in skip else max(sl["h2d"][0].elapsed_time(sl["h2d"][1]) for sl in slots)   # the last two uploads
    if prof is not None:
        import pstats
        prof.disable()
        pstats.Stats(prof, stream=sys.stderr).sort_stats("tottime").print_stats(16)
    if world > 1:
        tt = torch.tensor([dt], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
    return {"value": V * n / dt, "unit": UNIT, "ms_per_step": 1e3 * dt / n, "steps": n,
            "h2d_bytes_per_step": int(cam_host.numel() * 4 + gt_host.numel()), "d2h_bytes_per_step": 4 * L,
            "h2d_ms_per_step": round(float(h2d_ms), 4),
            "api": ("ShardedViewRasterizer.forward / .backward on this rank's parameter slice" if sr is not None else
                    "GaussianRasterizer(settings)(...) under autograd") +
                   " + l1_loss (mvs_gaussian_splatting_b200.loss) per rendered view + backward; the batch's camera blocks and "
                   "this rank's ground-truth images (8 bits per channel, as datasets store them; float32 on the device) come "
                   "from pinned host memory every step (prefetched one step ahead on a copy stream), the losses are copied back to pinned memory every step; Gaussian parameters resident in "
                   "HBM (model state)"}


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
    W, H, D = args.width, args.height, args.sh_degree
    P = args.points if args.points is not None else 6_000_000
    M = (D + 1) ** 2 if D == 0 else 16
    cam = S.front_camera(W, H)
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
