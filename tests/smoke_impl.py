"""smoke(): one small invocation of the hot path on cuda:0, checked against the oracle (C restatement)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def run():
    import __graft_entry__ as graft
    import oracle as O
    pkg = graft.load_package()
    S, B = pkg.scenes, pkg.binding
    ctx = B.Context(0)
    V, N, F = S.displaced_sphere(3, seed=5)
    sd = S.SceneDesc()
    sd.add_mesh(V, N, F)
    sd.add_rectangle(S.translate(0, -1.3, 0) @ S.rotate((1, 0, 0), -90) @ S.scale(2.0), emitter=None)
    sd.add_point_light((0, 5, 0), (1, 1, 1))
    rays = S.chord_rays(8192, radius=1.6, seed=11, mint=S.EPSILON)
    sc = sd.upload(ctx, B.BUILDER_PLOC)
    hits = sc.ray_intersect(rays)
    anyh = sc.ray_intersect_any(rays)
    osc = O.OracleScene(sd)
    t, u, v, prim = osc.closest(rays)
    ref_hit = np.isfinite(t)
    agree = np.where(ref_hit, hits["prim"] == prim, ~np.isfinite(hits["t"])).mean()
    assert agree >= 0.9995, "closest-hit agreement with the oracle %.5f" % agree
    m = ref_hit & (hits["prim"] == prim)
    assert np.abs(hits["t"][m] - t[m]).max() <= 1e-5 * 3.0
    assert (anyh == osc.any(rays)).mean() >= 0.9995
    print("smoke: %d rays, hit rate %.3f, primitive agreement %.5f, build %s" % (rays.shape[0], ref_hit.mean(), agree, sc.build_stats()))
    sc.close()
    try:
        import smoke_render
        smoke_render.run(pkg, ctx)
    except ImportError:
        pass
    ctx.close()
