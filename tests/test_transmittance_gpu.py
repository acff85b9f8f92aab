"""mtsb_eval_transmittance (the batched Scene::evalTransmittance, src/librender/scene.cpp:650-713) against the
unmodified reference on 6000 segments of the fog scene (tests/golden/transmittance_fog.npz, made by
tests/golden/make_golden_transmittance.py): three media, index-matched boundaries, a glass ball and opaque shapes in the way,
segments that name the wrong medium or none, with and without a bound on the number of boundary interactions."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("key,max_interactions", [("T", -1), ("T_max1", 1)])
def test_eval_transmittance_matches_reference(pkg, ctx, golden, key, max_interactions):
    g = golden("transmittance_fog.npz")
    sd = pkg.scenes.fog_scene(48, 32, spp=4, haze=True)
    sc = sd.upload(ctx)
    T = sc.eval_transmittance(g["p1"], g["p2"], g["medium"], p1_on_surface=(g["flags"] & 1) != 0, p2_on_surface=(g["flags"] & 2) != 0,
                              max_interactions=max_interactions)
    ref = g[key]
    # blocked / free segments agree exactly; the rest to float accuracy of exp(-sigma_t * d) summed over the media crossed
    zero_ref, zero = ref.max(axis=1) == 0, T.max(axis=1) == 0
    assert (zero_ref == zero).mean() >= 0.9995, (zero_ref != zero).sum()        # a segment grazing a boundary may flip
    both = ~zero_ref & ~zero
    assert both.mean() > 0.7
    assert np.abs(T[both] - ref[both]).max() <= 2e-6 + 1e-5 * ref[both].max()
    assert (T[ref.min(axis=1) == 1] == 1).all()
    if max_interactions == 1:
        assert (g["T"] != ref).any()          # the bound changes some answers in the reference, and the device follows
    sc.close()


def test_eval_transmittance_edge_cases(pkg, ctx):
    sd = pkg.scenes.fog_scene(48, 32, spp=4, haze=True)
    sc = sd.upload(ctx)
    p = np.array([[0.0, 1.5, 0.0]], np.float32)
    assert (sc.eval_transmittance(p, p, [2]) == 1).all()                          # p1 == p2: while (remaining > 0) never runs
    assert sc.eval_transmittance(np.zeros((0, 3)), np.zeros((0, 3)), []).shape == (0, 3)
    with pytest.raises(pkg.binding.MtsbError, match="names medium"):
        sc.eval_transmittance(p, p + 1, [7])
    # without media on the way the transmittance is 1 or 0: the answer of the any-hit query
    q = np.array([[0.0, 3.5, 0.0]], np.float32)
    assert (sc.eval_transmittance(p, q, [-1]) == 1).all()
    assert (sc.eval_transmittance(p, np.array([[0.0, -1.0, 0.0]], np.float32), [-1]) == 0).all()      # through the floor
    sc.close()
