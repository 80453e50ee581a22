"""Multi-view sharding + gradient all-reduce on CPU: world_size 2, gloo backend.  The per-view
fwd+bwd is injected (the CPU oracle stands in for the CUDA rasterizer, as the checker), so this
covers the host logic of mvs_gaussian_splatting_b200.multiview exactly as the N>1 bench runs it."""
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


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _scene():
    import helpers as Hh
    from mvs_gaussian_splatting_b200 import synthetic as S
    W, H, P, M = 96, 64, 600, 4
    cams = [S.front_camera(W, H)] + S.ring_cameras(2, W, H, radius=0.4)
    g = S.make_gaussians(P, cams[0], sh_coeffs=M, seed=3)
    dLs = [np.random.default_rng(10 + k).normal(size=(3, H, W)).astype(np.float32) for k in range(len(cams))]
    return Hh, cams, g, dLs, P, M


def _oracle_view(Hh, O, cam, g, dL):
    r = Hh.run_oracle(O, cam, g, [0, 0, 0], 1, dL_dpix=dL, threads=1)
    return r["grads"], r["radii"]


def _worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, "oracle"), HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import gs_oracle as O
    from mvs_gaussian_splatting_b200 import GradientArena
    from mvs_gaussian_splatting_b200.multiview import ViewStats, multiview_step, partition_views
    Hh, cams, g, dLs, P, M = _scene()
    arena, accum, stats = GradientArena(P, M, "cpu"), GradientArena(P, M, "cpu"), ViewStats(P, "cpu")

    def render_view(v, ar):
        grads, radii = _oracle_view(Hh, O, cams[v], g, dLs[v])
        ar.means3D.copy_(torch.from_numpy(grads["means3D"]))
        ar.shs.copy_(torch.from_numpy(grads["shs"]))
        ar.opacities.copy_(torch.from_numpy(grads["opacities"]))
        ar.scales.copy_(torch.from_numpy(grads["scales"]))
        ar.rotations.copy_(torch.from_numpy(grads["rotations"]))
        return torch.from_numpy(radii)

    mine = partition_views(len(cams), rank, world)
    out = multiview_step(mine, render_view, arena, accum, stats)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), flat=out.flat.numpy(), max_radii=stats.max_radii.numpy(),
             visible=stats.visible_count.numpy(), mine=np.array(mine))
    dist.barrier()
    dist.destroy_process_group()


def test_partition_views_covers_every_view_once():
    from mvs_gaussian_splatting_b200.multiview import partition_views
    for n in (0, 1, 7, 8, 13):
        for world in (1, 2, 4, 8):
            parts = [partition_views(n, r, world) for r in range(world)]
            assert sorted(v for p in parts for v in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        partition_views(4, 2, 2)


@pytest.mark.timeout(300)
def test_two_rank_step_equals_single_process_sum(tmp_path, oracle):
    world = 2
    port = _free_port()
    mp.start_processes(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True, start_method="spawn")
    r0, r1 = (np.load(tmp_path / f"rank{r}.npz") for r in range(world))
    assert list(r0["mine"]) == [0, 2] and list(r1["mine"]) == [1]
    np.testing.assert_array_equal(r0["flat"], r1["flat"])            # identical on every rank
    np.testing.assert_array_equal(r0["max_radii"], r1["max_radii"])

    Hh, cams, g, dLs, P, M = _scene()
    from mvs_gaussian_splatting_b200 import GradientArena
    want = GradientArena(P, M, "cpu")
    max_r, vis = np.zeros(P, np.int32), np.zeros(P, np.int32)
    for cam, dL in zip(cams, dLs):
        grads, radii = _oracle_view(Hh, oracle, cam, g, dL)
        for k in ("means3D", "shs", "opacities", "scales", "rotations"):
            getattr(want, k).add_(torch.from_numpy(grads[k]))
        max_r = np.maximum(max_r, radii)
        vis += radii > 0
    np.testing.assert_allclose(r0["flat"], want.flat.numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_array_equal(r0["max_radii"], max_r)
    np.testing.assert_array_equal(r0["visible"], vis)
