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
