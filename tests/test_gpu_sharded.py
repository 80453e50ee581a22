"""Multi-GPU parity of the Gaussian-sharded step (needs >= 2 CUDA devices; skipped otherwise).

Launches tools/sharded_check.py under torchrun on 2 ranks for both transports: image / radii / sorted
key lists of every rank's view bit-exact against the single-GPU path, slice gradients = per-view
single-GPU gradients summed over the views (5e-5 relative L1; north_star bar 1e-4)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
@pytest.mark.parametrize("exchange", ["peer", "nccl"])
def test_sharded_step_matches_single_gpu(cuda, exchange):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, GSR_EXCHANGE=exchange)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "sharded_check.py"), "60000", "640", "360"],
                       capture_output=True, text=True, env=env, timeout=500)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert f"sharded_check[{exchange}] world=2" in r.stdout and ": OK" in r.stdout
