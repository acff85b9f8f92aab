"""GPU tests of the `ldsampler` (csrc/shading.cuh ld_next1D / ld_next2D, mtsb_render_set_ldsampler).

The reference's LowDiscrepancySampler scrambles and shuffles its per-pixel point sets from a sequential Random in scheduler order
(src/samplers/ldsampler.cpp:147-206; its documentation says two runs differ), so there is no realisation to share.  What is pinned:
  * the point-set construction: the device's draws equal tests/ld_restate.py bit for bit, whose radicalInverse2Single /
    sobol2Single are pinned on the reference's own functions (tests/golden/qmc.npz, CPU test);
  * the properties the reference's sets have: every 2D set of the first `dimension` is a (0,2)-net in base 2 (one point in each
    elementary interval of area 1/N), every 1D set has one point per 1/N stratum, later draws are independent;
  * the images: mean and error of a 16-spp Cornell box against the reference's own ldsampler renders (four realisations) and its
    converged image (tests/golden/li_ldsampler_cornell.npz)."""
import numpy as np
import pytest

import ld_restate as L

pytestmark = pytest.mark.gpu


def _render(pkg, ctx, spp=16, dimension=4, seed=3, res=64):
    sd = pkg.scenes.cornell_box(res, res, spp=spp, max_depth=5, sampler="ldsampler", seed=seed, rfilter="box")
    sd.sampler["dimension"] = dimension
    sc = sd.upload(ctx)
    return sd, sc, sd.render(sc)


def _is_02_net(pts, log2n):
    n = 1 << log2n
    for a in range(log2n + 1):
        b = log2n - a
        cell = np.floor(pts[:, 0] * (1 << a)).astype(np.int64) * (1 << b) + np.floor(pts[:, 1] * (1 << b)).astype(np.int64)
        if np.unique(cell).size != n:
            return False
    return True


@pytest.mark.parametrize("spp,dimension", [(64, 4), (16, 2), (5, 4)])
def test_ldsampler_point_sets(pkg, ctx, spp, dimension):
    sd, sc, rd = _render(pkg, ctx, spp=spp, dimension=dimension, seed=9)
    log2n = int(np.ceil(np.log2(spp)))
    n, draws = 1 << log2n, dimension + 2
    seen = []
    for (px, py) in [(0, 0), (17, 5), (63, 63)]:
        got = rd.sampler_dump(px, py, 0, n, draws)
        want = L.dump(px, py, 0, n, draws, seed=9, dimension=dimension, log2n=log2n)
        assert (got.view(np.uint32) == want.view(np.uint32)).all()
        assert got.min() >= 0.0 and got.max() < 1.0
        for k in range(dimension):
            assert _is_02_net(got[:, k, 0:2], log2n), (px, py, k)
            assert np.unique(np.floor(got[:, k, 2] * n).astype(np.int64)).size == n, (px, py, k)
        seen.append(got)
    if n >= 64:
        # beyond `dimension` the draws are independent: 64 of them do not fill 64 strata
        assert np.unique(np.floor(seen[0][:, dimension, 2] * n).astype(np.int64)).size < n
        assert not _is_02_net(seen[0][:, dimension, 0:2], log2n)
    # different pixels, different scrambles and orders
    assert not np.array_equal(seen[0], seen[1]) and not np.array_equal(seen[1], seen[2])
    sc.close()


def _trimmed_rel_mse(img, hi, drop=0.02):
    e = np.sort(((img - hi) ** 2 / (hi ** 2 + 1e-3)).mean(axis=2).ravel())
    return float(e[: int(e.size * (1.0 - drop))].mean())


def test_ldsampler_image_statistics_match_the_reference(pkg, ctx, golden):
    g = golden("li_ldsampler_cornell.npz")
    hi = g["ref_hi"]
    ref_ld = [_trimmed_rel_mse(i, hi) for i in g["ref_ld"]]
    ref_ind = [_trimmed_rel_mse(i, hi) for i in g["ref_ind"]]
    errs, means = [], []
    for seed in (1, 2, 3, 4):
        sd, sc, rd = _render(pkg, ctx, spp=16, seed=seed)
        rd.render_pass(0, 16, 16)
        film = rd.film_to_host()
        b = rd.film.border
        img = film[b:film.shape[0] - b, b:film.shape[1] - b]
        img = img[..., :3] / img[..., 4:5]
        errs.append(_trimmed_rel_mse(img, hi)); means.append(float(img.mean()))
        sc.close()
    # unbiased: the mean over the image agrees with the converged reference
    assert abs(np.mean(means) - hi.mean()) <= 0.01 * hi.mean(), (means, hi.mean())
    # as good as the reference's own ldsampler (four realisations each), and better than its independent sampler
    assert np.mean(errs) <= 1.25 * np.mean(ref_ld), (errs, ref_ld)
    assert np.mean(errs) >= 0.6 * np.mean(ref_ld), (errs, ref_ld)
    assert np.mean(errs) < 0.9 * np.mean(ref_ind), (errs, ref_ind)


def test_ldsampler_through_the_wavefront_and_the_tail_agree(pkg, ctx):
    """Same sampler state through the wavefront stages and the late-bounce kernel: per-sample radiance identical."""
    sd, sc, rd = _render(pkg, ctx, spp=8, seed=5, res=32)
    ys, xs, ss = np.meshgrid(np.arange(32), np.arange(32), np.arange(8), indexing="ij")
    samples = np.stack([xs.ravel(), ys.ravel(), ss.ravel(), np.zeros(xs.size)], axis=1).astype(np.uint32)
    rd.set_tail(0)
    a = rd.render_samples(samples, 8)
    rd.set_tail(1 << 30)
    b = rd.render_samples(samples, 8)
    ok = (np.abs(a - b) <= 1e-5 + 1e-4 * np.abs(a)).all(axis=1)
    assert ok.mean() >= 0.999, ok.mean()
    sc.close()
