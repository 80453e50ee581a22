"""Multi-GPU parity of the Gaussian-sharded step against the single-GPU path (run under torchrun):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29533 tools/sharded_check.py [P W H]

Every rank renders ITS view with the ordinary single-GPU entry points on the full parameter set
(the reference-shaped path, itself checked against the oracle by tests/test_gpu_*.py) and with
ShardedViewRasterizer on its slice.  Checked: image / radii / sorted key list bit-exact (same
kernels, same geometry rows), slice gradients = sum over views of the single-GPU per-view
gradients within 5e-5 relative L1 (north_star bar: 1e-4; blend-backward atomics and the sum over
views commute only up to fp32 rounding).
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


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    W = int(sys.argv[2]) if len(sys.argv) > 2 else 800
    H = int(sys.argv[3]) if len(sys.argv) > 3 else 600
    D, M = 3, 16
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    sys.path.insert(0, ROOT)
    import bench  # camera ring of the benchmark
    g = S.make_gaussians(P, bench.bench_camera(0, W, H), sh_coeffs=M, seed=0)
    t = {k: torch.from_numpy(v).to(dev) for k, v in g.items()}
    cams = []
    for v in range(world):
        kw = S.settings_kwargs(bench.bench_camera(v, W, H), [0.1 * v, 0.2, 0.3], D)
        for k in ("bg", "viewmatrix", "projmatrix", "campos"):
            kw[k] = torch.from_numpy(np.ascontiguousarray(kw[k])).to(dev)
        cams.append(GaussianRasterizationSettings(**kw))
    gen = torch.Generator(device="cpu").manual_seed(7)
    dLs = [torch.randn(3, H, W, generator=gen).to(dev) for _ in range(world)]

    # ---- single-GPU reference: every view on this rank, gradients summed in view order ----
    want = None
    for v in range(world):
        color, radii, ctx, inp = rasterize_forward(t["means3D"], t["shs"], None, t["opacities"], t["scales"],
                                                   t["rotations"], None, cams[v])
        gr = rasterize_backward(ctx, inp, cams[v], dLs[v])
        torch.cuda.synchronize()
        cur = dict(means3D=gr[0], means2D=gr[1], shs=gr[2], opacities=gr[4], scales=gr[5], rotations=gr[6])
        if v == rank:
            ref_color, ref_radii = color.clone(), radii.clone()
            ref_keys = ctx.array("keys")[:ctx.num_rendered].clone()
            ref_pl = ctx.array("point_list")[:ctx.num_rendered].clone()
        if want is None:
            want = {k: x.clone() for k, x in cur.items() if k != "means2D"}
            m2d = [cur["means2D"].clone()]
        else:
            for k in want:
                want[k] += cur[k]
            m2d.append(cur["means2D"].clone())

    # ---- sharded ----
    exchange = os.environ.get("GSR_EXCHANGE", "peer")
    sr = ShardedViewRasterizer(P, M, W, H, dev, exchange=exchange)
    rows = sr.plan.rows(rank)
    sl = {k: t[k][rows] for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    for it in range(2):  # second pass exercises buffer reuse + capacity speculation
        color, radii = sr.forward(sl, cams)
        grads, gm2d = sr.backward(dLs[rank])
    torch.cuda.synchronize()
    ctx = sr._state[0]
    ok = True

    def check(name, cond):
        nonlocal ok
        if not cond:
            ok = False
            print(f"[rank {rank}] FAIL {name}", flush=True)

    check("image bit-exact", torch.equal(color, ref_color))
    check("radii bit-exact", torch.equal(radii, ref_radii))
    check("num_rendered", ctx.num_rendered == ref_keys.numel())
    check("keys bit-exact", torch.equal(ctx.array("keys")[:ctx.num_rendered], ref_keys))
    check("point_list bit-exact", torch.equal(ctx.array("point_list")[:ctx.num_rendered], ref_pl))
    worst = 0.0
    for k in want:
        a, b = grads[k].double(), want[k][rows].double()
        rel = float((a - b).abs().sum() / b.abs().sum().clamp_min(1e-30))
        worst = max(worst, rel)
        check(f"grad {k} rel_l1 {rel:.2e}", rel <= 5e-5)
    for v in range(world):
        a, b = gm2d[v].double(), m2d[v][rows].double()
        rel = float((a - b).abs().sum() / b.abs().sum().clamp_min(1e-30))
        check(f"means2D grad of view {v} rel_l1 {rel:.2e}", rel <= 5e-5)
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print(f"sharded_check[{exchange}] world={world} P={P} {W}x{H}: {'OK' if flag.item() == 0 else 'FAILED'} "
              f"(worst gradient rel_l1 {worst:.2e})", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 0 else 1)


if __name__ == "__main__":
    main()
