"""Host-side parameter derivations the reference performs in BSDF constructors (numpy/float64, run once per BSDF).

fresnel_diffuse_reflectance: fresnelDiffuseReflectance(eta, false) (src/libcore/util.cpp:816-870): for the non-fast
path the reference integrates fresnelDielectricExt(cosTheta, eta) * 2 cosTheta over [0,1] with a 128-point
Gauss-Lobatto rule; a dense composite Simpson rule on the same integrand agrees to < 1e-6.
"""
import math


def fresnel_dielectric_ext(cos_theta_i, eta):
    """fresnelDielectricExt (src/libcore/util.cpp:653-683), unpolarised reflectance only."""
    if eta == 1.0:
        return 0.0
    scale = 1.0 / eta if cos_theta_i > 0 else eta
    cos_t_sqr = 1.0 - (1.0 - cos_theta_i * cos_theta_i) * scale * scale
    if cos_t_sqr <= 0.0:
        return 1.0
    ci = abs(cos_theta_i)
    ct = math.sqrt(cos_t_sqr)
    rs = (ci - eta * ct) / (ci + eta * ct)
    rp = (eta * ci - ct) / (eta * ci + ct)
    return 0.5 * (rs * rs + rp * rp)


def fresnel_diffuse_reflectance(eta, n=4096):
    h = 1.0 / n
    s = 0.0
    for i in range(n + 1):
        x = i * h
        w = 1 if i in (0, n) else (4 if i % 2 else 2)
        s += w * fresnel_dielectric_ext(x, eta) * 2.0 * x
    return s * h / 3.0


# ---------------------------------------------------------------------------------------------- rough transmittance
# `roughplastic` reads precomputed rough-transmittance tables from the renderer's data directory
# (data/microfacet/{beckmann,ggx}.dat, loaded by RoughTransmittance, src/bsdfs/rtrans.h:80-150) and reduces the 3D table
# (eta x alpha x theta) to the 1D slice of its own eta and alpha with cubic interpolation (setEta :277-327, setAlpha
# :335-370; evalCubicInterp{1,2,3}D, src/libcore/spline.cpp:23-60, 236-304, 379-460).  Restated here in float32 numpy for
# hosts that drive the C ABI directly (the Mitsuba plugin does the same with the reference's own spline functions).
def _cubic_weights(p, size):
    """Knot index and the four node weights of evalCubicInterp{2,3}D for positions p in [0, 1] along an axis of `size`
    uniformly spaced knots (spline.cpp:241-287)."""
    import numpy as np
    f = np.float32
    p = np.atleast_1d(np.asarray(p, np.float32))
    t = (p * f(size - 1)) / f(1.0)
    knot = np.minimum(t.astype(np.int64), size - 2)
    t = (t - knot.astype(np.float32)).astype(np.float32)
    t2 = t * t
    t3 = t2 * t
    w = np.zeros((p.shape[0], 4), np.float32)
    w[:, 1] = f(2) * t3 - f(3) * t2 + f(1)
    w[:, 2] = f(-2) * t3 + f(3) * t2
    d0 = t3 - f(2) * t2 + t
    d1 = t3 - t2
    lo = knot > 0
    w[:, 2] += np.where(lo, f(0.5) * d0, d0)
    w[:, 0] -= np.where(lo, f(0.5) * d0, f(0))
    w[:, 1] -= np.where(lo, f(0), d0)
    hi = knot + 2 < size
    w[:, 3] += np.where(hi, f(0.5) * d1, f(0))
    w[:, 1] -= np.where(hi, f(0.5) * d1, d1)
    w[:, 2] += np.where(hi, f(0), d1)
    return knot, w


def _interp_axis(values, axis, p):
    """Cubic interpolation of `values` along `axis` at positions p (one per output entry along that axis)."""
    import numpy as np
    size = values.shape[axis]
    knot, w = _cubic_weights(p, size)
    out = None
    for k in range(4):
        idx = np.clip(knot + (k - 1), 0, size - 1)
        term = np.take(values, idx, axis=axis) * np.expand_dims(w[:, k], tuple(a for a in range(values.ndim) if a != axis))
        out = term if out is None else out + term
    return out.astype(np.float32)


def load_rough_transmittance(path):
    """-> dict(trans[2*eta, alpha, theta], diff[2*eta, alpha], eta_min, eta_max, alpha_min, alpha_max)"""
    import numpy as np
    import struct
    raw = open(path, "rb").read()
    if raw[:17] != b"MTS_TRANSMITTANCE":
        raise ValueError("%s is not a rough transmittance table" % path)
    n_eta, n_alpha, n_theta = struct.unpack("<QQQ", raw[17:41])
    eta_min, eta_max, alpha_min, alpha_max = struct.unpack("<ffff", raw[41:57])
    data = np.frombuffer(raw, "<f4", offset=57).reshape(2 * n_eta, n_alpha, n_theta + 1)
    return dict(trans=np.ascontiguousarray(data[..., :n_theta]), diff=np.ascontiguousarray(data[..., n_theta]),
                n_eta=n_eta, eta_min=np.float32(eta_min), eta_max=np.float32(eta_max), alpha_min=np.float32(alpha_min), alpha_max=np.float32(alpha_max))


def rough_transmittance_slices(table, eta, alpha):
    """What RoughPlastic::configure keeps (roughplastic.cpp:283-303): the external table after setEta(eta), setAlpha(alpha)
    (n_theta values over cosTheta^(1/4)) and Fdr = 1 - evalDiffuse(alpha) of the internal table after setEta(1 / eta)."""
    import numpy as np
    f = np.float32

    def set_eta(e):
        e = f(e)
        trans, diff = table["trans"], table["diff"]
        n = table["n_eta"]
        if e < 1:
            trans, diff = trans[n:], diff[n:]
            e = f(1) / e
        else:
            trans, diff = trans[:n], diff[:n]
        e = max(e, table["eta_min"])
        warped = np.power((e - table["eta_min"]) / (table["eta_max"] - table["eta_min"]), f(0.25), dtype=np.float32)
        n_alpha, n_theta = trans.shape[1], trans.shape[2]
        t = _interp_axis(trans, 0, [warped])[0]                                   # -> [alpha, theta]
        # the knots of the other two axes are hit at i * (1 / (n - 1)) in float32, as setEta does
        t = _interp_axis(t, 0, np.arange(n_alpha, dtype=np.float32) * (f(1) / f(n_alpha - 1)))
        t = _interp_axis(t, 1, np.arange(n_theta, dtype=np.float32) * (f(1) / f(n_theta - 1)))
        d = _interp_axis(diff, 0, [warped])[0]
        d = _interp_axis(d, 0, np.arange(n_alpha, dtype=np.float32) * (f(1) / f(n_alpha - 1)))
        return t, d

    warped_alpha = np.power((f(alpha) - table["alpha_min"]) / (table["alpha_max"] - table["alpha_min"]), f(0.25), dtype=np.float32)
    ext_t, _ = set_eta(eta)
    n_theta = ext_t.shape[1]
    ext = _interp_axis(ext_t, 0, [warped_alpha])[0]
    ext = _interp_axis(ext, 0, np.arange(n_theta, dtype=np.float32) * (f(1) / f(n_theta - 1)))
    _, int_d = set_eta(f(1) / f(eta))
    diff_int = float(np.clip(_interp_axis(int_d, 0, [warped_alpha])[0], 0.0, 1.0))
    return ext.astype(np.float32), 1.0 - diff_int
