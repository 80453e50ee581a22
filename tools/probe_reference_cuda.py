"""Run-time probe: is the reference's CUDA rasterizer — the un-vendored
submodule `diff_gaussian_rasterization` (/root/reference/.gitmodules:4-6), or `gsplat` — installed on
THIS machine outside this repository?  If it is, it is the primary oracle and the timed CUDA baseline;
if not, parity stays pinned to the CPU oracle and the record says so.

The repo ships a shim package of the same import name (the drop-in boundary), so the probe removes the
repo root from sys.path before looking, and rejects any hit that resolves inside the repo.

    python tools/probe_reference_cuda.py            # prints one JSON object
"""
from __future__ import annotations

import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def probe() -> dict:
    saved = list(sys.path)
    try:
        sys.path[:] = [p for p in sys.path if os.path.abspath(p or os.getcwd()) != ROOT]
        for m in [k for k in sys.modules if k.split(".")[0] in ("diff_gaussian_rasterization", "gsplat")]:
            del sys.modules[m]
        importlib.invalidate_caches()
        out = {}
        for name in ("diff_gaussian_rasterization", "gsplat", "diff_surfel_rasterization", "simple_knn"):
            try:
                spec = importlib.util.find_spec(name)
            except (ImportError, ValueError):
                spec = None
            origin = getattr(spec, "origin", None) if spec else None
            locs = list(getattr(spec, "submodule_search_locations", None) or []) if spec else []
            where = origin or (locs[0] if locs else None)
            if where and os.path.abspath(where).startswith(ROOT + os.sep):
                where = None   # this repo's own shim
            out[name] = where
        return out
    finally:
        sys.path[:] = saved


def summary() -> dict | str:
    found = {k: v for k, v in probe().items() if v}
    cuda_ref = {k: v for k, v in found.items() if k in ("diff_gaussian_rasterization", "gsplat")}
    return cuda_ref if cuda_ref else "not installed"


if __name__ == "__main__":
    import platform
    print(json.dumps({"reference_cuda": summary(), "all": probe(), "host": platform.node(),
                      "python": sys.version.split()[0]}))
