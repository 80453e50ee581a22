"""Parity of the Gaussian-sharded multi-view step (a V-view batch per step, rank r renders views r, r+N, ...).

tools/sharded_check.py compares it with the single-GPU path: image / radii / sorted key lists of every local
view bit-exact, slice gradients = per-view single-GPU gradients summed over the V views (5e-5 relative L1;
north_star bar 1e-4), per-view dL/dmeans2D, the slice's per-view radii and the densification statistics.

  * world 1 (one process, "local" transport, 3 views): runs on ANY GPU box — the multi-view kernels, the view
    plan and the statistics are covered even when the lease has a single device;
  * world 2 with 4 views (two per rank), both transports; world 4 with 8 views: skipped below 2 / 4 devices.
"""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHECK = os.path.join(ROOT, "tools", "sharded_check.py")


@pytest.mark.timeout(600)
@pytest.mark.parametrize("views", [1, 3])
def test_local_multi_view_step_matches_single_view_path(cuda, views):
    r = subprocess.run([sys.executable, CHECK, "60000", "640", "360", str(views)], capture_output=True, text=True,
                       env={k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")},
                       timeout=500)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert f"sharded_check[local] world=1 views={views}" in r.stdout and ": OK" in r.stdout


def _torchrun(world, exchange, views):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, GSR_EXCHANGE=exchange)
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                           "--master-addr", "127.0.0.1", "--master-port", str(port), CHECK, "60000", "640", "360",
                           str(views)], capture_output=True, text=True, env=env, timeout=500)


@pytest.mark.timeout(600)
@pytest.mark.parametrize("exchange,views", [("peer", 2), ("peer", 4), ("nccl", 4)])
def test_sharded_step_matches_single_gpu(cuda, exchange, views):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    r = _torchrun(2, exchange, views)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert f"sharded_check[{exchange}] world=2 views={views}" in r.stdout and ": OK" in r.stdout


@pytest.mark.timeout(600)
def test_sharded_step_eight_views_on_four_gpus(cuda):
    import torch
    if torch.cuda.device_count() < 4:
        pytest.skip("needs 4 CUDA devices")
    r = _torchrun(4, "peer", 8)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert "sharded_check[peer] world=4 views=8" in r.stdout and ": OK" in r.stdout


@pytest.mark.timeout(300)
def test_second_device_in_one_process(cuda, oracle):
    """One process renders on cuda:0 and then on cuda:1 (the dynamic shared-memory opt-in of the radix passes
    and of the multi-view backward is per device)."""
    import numpy as np
    import torch
    import helpers as Hh
    from mvs_gaussian_splatting_b200 import GaussianRasterizer
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    cam, g = Hh.make_case(5000, 320, 200, seed=2)
    ref = Hh.run_oracle(oracle, cam, g, [0, 0, 0], 3)
    for idx in (0, 1):
        dev = torch.device("cuda", idx)
        rs = Hh.settings_for(cam, [0, 0, 0], 3, dev)
        t = Hh.to_dev(g, dev, requires_grad=True)
        m2 = torch.zeros_like(t["means3D"], requires_grad=True)
        color, radii = GaussianRasterizer(rs)(means3D=t["means3D"], means2D=m2, shs=t["shs"], opacities=t["opacities"],
                                              scales=t["scales"], rotations=t["rotations"])
        color.sum().backward()
        torch.cuda.synchronize(dev)
        np.testing.assert_array_equal(radii.cpu().numpy(), ref["radii"])
        assert bool(torch.isfinite(t["means3D"].grad).all())
