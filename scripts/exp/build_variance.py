"""Where do the 700-800 ms of `build_other` (PLOC build of the config-2 mesh inside bench.py) come from?  The same build takes
14 ms in a fresh process.  Rebuilds the mesh with both builders after each step bench.py takes before it."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
import bench  # noqa: E402

pkg = g.load_package()
B, S = pkg.binding, pkg.scenes


def builds(ctx, sd, tag):
    for name, b in (("ploc", B.BUILDER_PLOC), ("lbvh", B.BUILDER_LBVH), ("ploc", B.BUILDER_PLOC)):
        t0 = time.perf_counter()
        sc = sd.upload(ctx, b)
        ctx.sync()
        wall = time.perf_counter() - t0
        print("%-28s %s build %.2f ms (upload+commit wall %.1f ms)" % (tag, name, sc.build_stats()["build_ms"], wall * 1e3), flush=True)
        sc.close()


def main():
    import torch
    ctx = B.Context(0)
    sd = bench.build_mesh(S)
    builds(ctx, sd, "fresh")
    sc = sd.upload(ctx, B.BUILDER_LBVH)
    n = 1 << 24
    rays = S.chord_rays(n, seed=1) if hasattr(S, "chord_rays") else None
    if rays is None:
        print("no chord_rays"); return
    rays_d = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).cuda()
    hits_d = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    builds(ctx, sd, "after torch allocations")
    sc.trace_device("closest", rays_d.data_ptr(), n, hits_d.data_ptr())
    torch.cuda.synchronize()
    builds(ctx, sd, "after a device batch (L2 set-aside on)")
    pin_r, pin_h = B.PinnedArray((n,), B.ray_dtype), B.PinnedArray((n,), B.hit_dtype)
    pin_r.array[:] = rays
    sc.ray_intersect(pin_r.array, out=pin_h.array)
    builds(ctx, sd, "after a host batch (pipe)")


main()
