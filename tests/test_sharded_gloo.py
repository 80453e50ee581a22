"""CPU tests of the Gaussian-sharded multi-view step's host logic (mvs_gaussian_splatting_b200/sharded.py):
the slice plan, the batched all-to-all over gloo with world_size 2, and — with the oracle's stage
functions standing in for the CUDA kernels — that "preprocess slices per camera -> exchange geometry
rows -> bin/blend per view -> exchange accumulator slices -> preprocess_backward per view, summed"
reproduces the single-process rasterizer exactly (image, radii) and its per-view gradient sum."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def test_view_plan_assigns_every_view_once():
    from mvs_gaussian_splatting_b200.sharded import ViewPlan
    for world in (1, 2, 4, 8):
        for V in (world, 2 * world, 8 if 8 % world == 0 else world):
            vp = ViewPlan(V, world)
            seen = sorted(v for r in range(world) for v in vp.local_views(r))
            assert seen == list(range(V))
            for r in range(world):
                assert all(vp.owner(v) == r for v in vp.local_views(r))
                assert [vp.local_index(v) for v in vp.local_views(r)] == list(range(V // world))
                order = vp.order(r)
                assert sorted(order) == list(range(V)) and order[0] == r
            # at every position of the order the ranks address DIFFERENT owners (staggered peer traffic)
            for k in range(V):
                assert len({vp.owner(vp.order(r)[k]) for r in range(world)}) == world
    for bad in ((3, 2), (0, 1), (4, 0)):
        with pytest.raises(ValueError):
            ViewPlan(*bad)


def test_shard_plan_tiles_the_index_range():
    from mvs_gaussian_splatting_b200.sharded import SLICE_ALIGN, ShardPlan
    for P in (0, 1, 127, 128, 1000, 100_000, 3_000_000):
        for world in (1, 2, 3, 8):
            plan = ShardPlan(P, world)
            assert sum(plan.sizes) == P
            pos = 0
            for r in range(world):
                assert plan.starts[r] == pos and plan.starts[r] % SLICE_ALIGN == 0 or plan.sizes[r] == 0
                assert plan.rows(r) == slice(plan.starts[r], plan.starts[r] + plan.sizes[r])
                assert plan.slot_stride(r) % SLICE_ALIGN == 0 and plan.slot_stride(r) >= plan.sizes[r]
                pos += plan.sizes[r]
            # all but the last non-empty slice are whole multiples of the alignment
            nonempty = [s for s in plan.sizes if s]
            assert all(s % SLICE_ALIGN == 0 for s in nonempty[:-1])
    with pytest.raises(ValueError):
        ShardPlan(-1, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _scene():
    sys.path.insert(0, HERE)
    import helpers as Hh
    from mvs_gaussian_splatting_b200 import synthetic as S
    W, H, P, M = 96, 64, 700, 4
    cam0 = S.front_camera(W, H)
    g = S.make_gaussians(P, cam0, sh_coeffs=M, seed=11)
    cams = [cam0]
    for eye in ([0.3, -0.2, 0.0], [-0.4, 0.1, 0.0], [0.1, 0.35, 0.0]):   # a 4-view batch: two views per rank at world 2
        R, t = S.look_at(eye, [0.0, 0.0, 10.0])
        cams.append(S.make_camera(R, t, np.radians(60.0), W, H))
    rng = np.random.default_rng(5)
    dLs = [rng.normal(size=(3, H, W)).astype(np.float32) for _ in cams]
    return Hh, cams, g, dLs, P, M, W, H


GEOM = ("depths", "means2D", "conic_opacity", "rgb", "tiles_touched", "radii", "rects")


def _pre(O, cam, g, rows, D):
    s = {k: g[k][rows] for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    return O.preprocess(s["means3D"], s["scales"], s["rotations"], s["opacities"], s["shs"], cam.world_view_transform,
                        cam.full_proj_transform, cam.camera_center, cam.width, cam.height, cam.tanfovx, cam.tanfovy, D)


def _worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, "oracle"), HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import gs_oracle as O
    from mvs_gaussian_splatting_b200.sharded import ShardPlan, ViewPlan, all_to_all_rows
    Hh, cams, g, dLs, P, M, W, H = _scene()
    D = 1
    V = len(cams)
    plan, views = ShardPlan(P, world), ViewPlan(V, world)
    rows = plan.rows(rank)
    bg = np.zeros(3, np.float32)
    # forward: my slice under every camera of the batch, then the geometry rows of view v travel to rank owner(v)
    mine = [_pre(O, cams[v], g, rows, D) for v in range(V)]
    mine_local = views.local_views(rank)
    assert mine_local == [rank + l * world for l in range(V // world)]
    state = []
    for l, v_mine in enumerate(mine_local):
        full, pairs = {}, []
        for k in GEOM:
            a0 = np.ascontiguousarray(mine[0][k])
            full[k] = torch.zeros((P,) + a0.shape[1:], dtype=torch.from_numpy(a0).dtype)
            # rank j renders view j + l*world as ITS local view l
            pairs.append(([torch.from_numpy(np.ascontiguousarray(mine[j + l * world][k])) for j in range(world)],
                          [full[k][plan.rows(j)] for j in range(world)]))
        all_to_all_rows(pairs)
        pre = {k: full[k].numpy() for k in GEOM}
        binned = O.bin_instances(pre, W, H)
        fwd = O.blend_forward(pre, binned, bg, W, H)
        acc = O.blend_backward(pre, binned, fwd, bg, dLs[v_mine], W, H)
        state.append((pre, fwd, acc))
    # backward: accumulator slices of every local view travel to the rank that owns the slice
    names = sorted(k for k, v in state[0][2].items() if isinstance(v, np.ndarray) and v.shape[:1] == (P,))
    got = {k: [None] * V for k in names}
    for l in range(len(mine_local)):
        pairs = []
        for k in names:
            t = torch.from_numpy(np.ascontiguousarray(state[l][2][k]))
            recv = [torch.zeros((plan.sizes[rank],) + t.shape[1:], dtype=t.dtype) for _ in range(world)]
            for j in range(world):
                got[k][j + l * world] = recv[j]
            pairs.append(([t[plan.rows(j)] for j in range(world)], recv))
        all_to_all_rows(pairs)
    total = None
    s = {k: g[k][rows] for k in ("means3D", "scales", "rotations", "shs")}
    for v in views.order(rank):
        assert views.owner(v) == v % world and views.local_index(v) == v // world
        gb = {k: got[k][v].numpy() for k in names}
        cam = cams[v]
        out = O.preprocess_backward(s["means3D"], s["scales"], s["rotations"], s["shs"], cam.world_view_transform,
                                    cam.full_proj_transform, cam.camera_center, W, H, cam.tanfovx, cam.tanfovy, D,
                                    mine[v], gb)
        out = dict(means3D=out["dL_dmeans3D"], shs=out["dL_dsh"], scales=out["dL_dscales"], rotations=out["dL_drots"],
                   opacities=gb["dL_dopacity"].reshape(-1, 1))
        out = {k: np.asarray(x, np.float64) for k, x in out.items()}
        total = out if total is None else {k: total[k] + out[k] for k in total}
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), color=np.stack([st[1]["color"] for st in state]),
             radii=np.stack([st[0]["radii"] for st in state]), **{f"g_{k}": v for k, v in total.items()})
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_sharded_step_equals_single_process(tmp_path, oracle):
    world = 2
    mp.start_processes(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True,
                       start_method="spawn")
    Hh, cams, g, dLs, P, M, W, H = _scene()
    from mvs_gaussian_splatting_b200.sharded import ShardPlan
    plan = ShardPlan(P, world)
    V = len(cams)
    refs = [Hh.run_oracle(oracle, cams[v], g, [0, 0, 0], 1, dL_dpix=dLs[v], threads=1) for v in range(V)]
    for r in range(world):
        out = np.load(tmp_path / f"rank{r}.npz")
        for l in range(V // world):                                        # rank r renders views r, r + world, ...
            np.testing.assert_array_equal(out["color"][l], refs[r + l * world]["color"])      # bit-exact
            np.testing.assert_array_equal(out["radii"][l], refs[r + l * world]["radii"])
        for k in ("means3D", "shs", "opacities", "scales", "rotations"):
            want = sum(np.asarray(refs[v]["grads"][k], np.float64) for v in range(V))[plan.rows(r)]
            got = out[f"g_{k}"].reshape(want.shape)
            np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-9, err_msg=k)
