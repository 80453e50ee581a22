"""Parity of the Gaussian-sharded multi-view step against the single-GPU path.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29533 tools/sharded_check.py [P W H [V]]        # N ranks, V views per step (default 2N)
    python tools/sharded_check.py [P W H [V]]                          # one process: world 1, "local" transport

Every rank renders ITS views with the ordinary single-GPU entry points on the full parameter set (the
reference-shaped path, itself checked against the oracle by tests/test_gpu_*.py) and with
ShardedViewRasterizer on its slice.  Checked: image / radii / sorted key list of every local view bit-exact
(same kernels, same geometry rows), slice gradients = sum over the V views of the single-GPU per-view gradients
within 5e-5 relative L1 (north_star bar: 1e-4; blend-backward atomics and the sum over views commute only up to
fp32 rounding), per-view dL/dmeans2D rows, and the densification statistics of train.py:130-131 accumulated over
the batch against the reference's own index expressions.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from mvs_gaussian_splatting_b200 import (GaussianRasterizationSettings, rasterize_backward,  # noqa: E402
                                         rasterize_forward)
from mvs_gaussian_splatting_b200 import synthetic as S  # noqa: E402
from mvs_gaussian_splatting_b200.sharded import ShardedViewRasterizer  # noqa: E402


def run_check(P, W, H, V, dev, rank, world, exchange):
    """Returns (ok, worst gradient rel_l1) on this rank; prints every failed item."""
    D, M = 3, 16
    cam_list = S.batch_cameras(V, W, H)
    g = S.make_gaussians(P, S.front_camera(W, H), sh_coeffs=M, seed=0)
    t = {k: torch.from_numpy(v).to(dev) for k, v in g.items()}
    cams = []
    for v in range(V):
        kw = S.settings_kwargs(cam_list[v], [0.1 * v, 0.2, 0.3], D)
        for k in ("bg", "viewmatrix", "projmatrix", "campos"):
            kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k])).to(dev)
        cams.append(GaussianRasterizationSettings(**kw))
    gen = torch.Generator(device="cpu").manual_seed(7)
    dLs = [torch.randn(3, H, W, generator=gen).to(dev) for _ in range(V)]

    sr = ShardedViewRasterizer(P, M, W, H, dev, exchange=exchange, n_views=V)
    mine = sr.local_views()
    rows = sr.plan.rows(rank)

    # ---- single-GPU reference: every view on this rank, gradients summed in view order ----
    want, m2d, ref, all_radii = None, [], {}, []
    for v in range(V):
        color, radii, ctx, inp = rasterize_forward(t["means3D"], t["shs"], None, t["opacities"], t["scales"],
                                                   t["rotations"], None, cams[v])
        gr = rasterize_backward(ctx, inp, cams[v], dLs[v])
        torch.cuda.synchronize()
        cur = dict(means3D=gr[0], means2D=gr[1], shs=gr[2], opacities=gr[4], scales=gr[5], rotations=gr[6])
        if v in mine:
            ref[v] = (color.clone(), radii.clone(), ctx.array("keys")[:ctx.num_rendered].clone(),
                      ctx.array("point_list")[:ctx.num_rendered].clone())
        all_radii.append(radii.clone())
        m2d.append(cur["means2D"].clone())
        if want is None:
            want = {k: x.clone() for k, x in cur.items() if k != "means2D"}
        else:
            for k in want:
                want[k] += cur[k]

    # ---- sharded ----
    sl = {k: t[k][rows] for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    Ps = rows.stop - rows.start
    stats = [torch.zeros(Ps, device=dev), torch.zeros(Ps, 1, device=dev), torch.zeros(Ps, 1, device=dev)]
    for it in range(2):  # second pass exercises buffer reuse + capacity speculation
        colors, radii = sr.forward(sl, cams)
        grads, gm2d = sr.backward(torch.stack([dLs[v] for v in mine]))
    sr.densification_stats(*stats)
    torch.cuda.synchronize()
    ok = True

    def check(name, cond):
        nonlocal ok
        if not cond:
            ok = False
            print(f"[rank {rank}] FAIL {name}", flush=True)

    for l, v in enumerate(mine):
        ctx = sr._state[0][l]
        ref_color, ref_radii, ref_keys, ref_pl = ref[v]
        check(f"view {v} image bit-exact", torch.equal(colors[l], ref_color))
        check(f"view {v} radii bit-exact", torch.equal(radii[l], ref_radii))
        check(f"view {v} num_rendered", ctx.num_rendered == ref_keys.numel())
        check(f"view {v} keys bit-exact", torch.equal(ctx.array("keys")[:ctx.num_rendered], ref_keys))
        check(f"view {v} point_list bit-exact", torch.equal(ctx.array("point_list")[:ctx.num_rendered], ref_pl))
    worst = 0.0
    for k in want:
        a, b = grads[k].double(), want[k][rows].double()
        rel = float((a - b).abs().sum() / b.abs().sum().clamp_min(1e-30))
        worst = max(worst, rel)
        check(f"grad {k} rel_l1 {rel:.2e}", rel <= 5e-5)
    for v in range(V):
        a, b = gm2d[v].double(), m2d[v][rows].double()
        rel = float((a - b).abs().sum() / b.abs().sum().clamp_min(1e-30))
        check(f"means2D grad of view {v} rel_l1 {rel:.2e}", rel <= 5e-5)
        check(f"slice radii of view {v}", torch.equal(sr.slice_radii()[v], all_radii[v][rows]))
    # densification statistics over the batch, the reference's statements (train.py:130-131, gaussian_model.py:775-777)
    max_r, acc, den = torch.zeros(Ps, device=dev), torch.zeros(Ps, 1, device=dev), torch.zeros(Ps, 1, device=dev)
    for v in range(V):
        r_v, g_v = all_radii[v][rows], m2d[v][rows]
        vis = r_v > 0
        max_r[vis] = torch.max(max_r[vis], r_v[vis])
        acc[vis] += torch.norm(g_v[vis, :2], dim=-1, keepdim=True)
        den[vis] += 1
    check("densification max_radii2D", torch.equal(stats[0], max_r))
    check("densification denom", torch.equal(stats[2], den))
    rel = float((stats[1] - acc).abs().sum() / acc.abs().sum().clamp_min(1e-30))
    check(f"densification xyz_gradient_accum rel_l1 {rel:.2e}", rel <= 5e-5)
    return ok, worst


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    W = int(sys.argv[2]) if len(sys.argv) > 2 else 800
    H = int(sys.argv[3]) if len(sys.argv) > 3 else 600
    distributed = "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1
    if distributed:
        rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    else:
        rank, world, local = 0, 1, 0
    V = int(sys.argv[4]) if len(sys.argv) > 4 else 2 * world
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if distributed:
        dist.init_process_group("nccl", device_id=dev)
    exchange = os.environ.get("GSR_EXCHANGE", "peer") if distributed else "local"
    ok, worst = run_check(P, W, H, V, dev, rank, world, exchange)
    flag = torch.tensor([0 if ok else 1], device=dev)
    if distributed:
        dist.all_reduce(flag)
    if rank == 0:
        print(f"sharded_check[{exchange}] world={world} views={V} P={P} {W}x{H}: "
              f"{'OK' if flag.item() == 0 else 'FAILED'} (worst gradient rel_l1 {worst:.2e})", flush=True)
    if distributed:
        dist.barrier()
        dist.destroy_process_group()
    sys.exit(0 if flag.item() == 0 else 1)


if __name__ == "__main__":
    main()
