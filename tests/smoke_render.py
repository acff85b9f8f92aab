"""Second half of smoke(): one tiny wavefront render on cuda:0 checked against the golden image the unmodified reference
produced for the same scene and the same random numbers (tests/golden/li_cornell.npz)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(pkg, ctx):
    S, B = pkg.scenes, pkg.binding
    g = np.load(os.path.join(ROOT, "tests", "golden", "li_cornell.npz"))
    ref, spp, seed = g["image"], int(g["spp"]), int(g["seed"])
    sd = S.cornell_box(32, 32, spp=spp, max_depth=5, sampler="ctr", seed=seed)
    sc = sd.upload(ctx)
    rd = B.Render(sc, sd.camera(), sd.film(), sd.integrator_desc(), seed=seed)
    rd.render_pass(0, spp, spp)
    film = rd.film_to_host()
    b = sd.film().border
    film = film[b:film.shape[0] - b, b:film.shape[1] - b]
    img = film[..., 0:3] / film[..., 4:5]
    ok = (np.abs(img - ref) <= 2e-4 + 2e-3 * np.abs(ref)).all(axis=2).mean()
    st = rd.stats()
    assert ok >= 0.97, "render agreement with the reference image %.4f" % ok
    print("smoke render: 32x32x%d spp, pixel agreement with the reference %.4f, avg path length %.2f, %d launches" % (
        spp, ok, st["depth_sum"] / st["paths"], st["launches"]))
    rd.close()
    sc.close()
