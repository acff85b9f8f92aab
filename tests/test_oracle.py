"""Pins the plain-C restatement (oracle/restate.c) against golden vectors produced by the UNMODIFIED reference
(tests/golden/make_golden.py -> oracle/_ref -> ShapeKDTree::rayIntersect).  CPU only."""
import numpy as np
import pytest

import helpers
import oracle as O


@pytest.mark.parametrize("kind", ["sphere", "cornell"])
def test_restatement_matches_reference_hits(pkg, golden, kind):
    gold = golden("hits_%s.npz" % kind)
    sd = helpers.desc_from_golden(pkg.scenes, gold, kind)
    osc = O.OracleScene(sd)
    t, u, v, prim = osc.closest(gold["rays"])
    # reference reports (shape, local prim); the restatement the kd-tree's global index
    offs = np.cumsum([0] + [s["F"].shape[0] if s["type"] == "mesh" else 1 for s in sd.shapes])
    ref_prim = np.where(gold["ref_shape"] >= 0, offs[np.maximum(gold["ref_shape"], 0)] + np.maximum(gold["ref_prim"], 0), -1)
    is_rect = np.array([s["type"] != "mesh" for s in sd.shapes])
    # Rectangle::fillIntersectionRecord leaves primIndex unset; compare by shape there
    rect_hit = (gold["ref_shape"] >= 0) & is_rect[np.maximum(gold["ref_shape"], 0)]
    ref_prim = np.where(rect_hit, offs[np.maximum(gold["ref_shape"], 0)], ref_prim)
    # the Cornell set contains 14 shadow rays that start ON the light rectangle and run inside its plane (|d.n| < 1e-5):
    # their self-intersection is decided by the last bit of -o.z / d.z in Rectangle::rayIntersect.  They are tagged by
    # that geometric criterion and left out; every other ray is held to the north-star bound (99.99 %)
    rays = gold["rays"]
    graze = (np.abs(rays[:, 1] - 1.98) < 1e-4) & (np.abs(rays[:, 5]) < 1e-5) if kind == "cornell" else np.zeros(len(rays), bool)
    st = helpers.compare_hits(rays, t, prim.view(np.int32), gold["ref_t"], ref_prim, min_agree=helpers.PRIM_AGREE, exclude=graze)
    assert st["miss_agree"] >= helpers.PRIM_AGREE and st["max_dt_unweighted"] <= helpers.T_REL
    a = osc.any(rays)
    assert (a == gold["ref_any"])[~graze].mean() >= helpers.PRIM_AGREE
    # barycentrics of the reference (its.uv = (b1, b2) on a mesh without texture coordinates) on the sphere set
    if kind == "sphere":
        m = np.isfinite(t) & (prim.view(np.int32) == ref_prim)
        f = gold["F"][np.maximum(gold["ref_prim"], 0)]
        e1, e2 = gold["V"][f[:, 1]] - gold["V"][f[:, 0]], gold["V"][f[:, 2]] - gold["V"][f[:, 0]]
        cos = np.abs((rays[:, 4:7] * gold["ref_ng"]).sum(axis=1))
        helpers.compare_barycentrics(rays[m], u[m], v[m], gold["ref_u"][m], gold["ref_v"][m], e1[m], e2[m], gold["ref_t"][m], cos=cos[m])


def test_restatement_aabb_is_enlarged_like_the_kdtree(pkg):
    S = pkg.scenes
    V, N, F = S.displaced_sphere(1, seed=3)
    sd = S.SceneDesc()
    sd.add_mesh(V, N, F)
    mn, mx = O.OracleScene(sd).aabb()
    tmn, tmx = V.min(0), V.max(0)
    lo = tmn - ((tmx - tmn) * np.float32(1e-3) + np.float32(1e-3))
    hi = tmx + ((tmx - lo) * np.float32(1e-3) + np.float32(1e-3))
    assert np.allclose(mn, lo, rtol=0, atol=1e-6) and np.allclose(mx, hi, rtol=0, atol=1e-6)
