"""GPU parity tests for the shading half of the hot path, through the C ABI: BSDF / emitter plugin math pointwise against
the reference's plugins, per-sample radiance against MIPathTracer::Li under the shared counter-based sampler, film
splatting against ImageBlock::put, and image-level unbiasedness against a live reference render."""
import os

import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu


def _pyramid(golden):
    """The MIP pyramid the reference builds from scenes.sky_image() (tests/golden/envmap_pyramid.npz, raydump `mipmap`)."""
    g = golden("envmap_pyramid.npz")
    return [g["level%d" % l].astype(np.float32) for l in range(len(g.files))]


def _tex_pyramids(golden):
    """The three MIP pyramids (repeat, clamp, mirror) the reference builds from scenes.texture_image() (raydump `mipmap3`)."""
    g = golden("texture_pyramids.npz")
    out = []
    for wrap in ("repeat", "clamp", "mirror"):
        n = sum(1 for k in g.files if k.startswith(wrap + "_level"))
        out.append([g["%s_level%d" % (wrap, l)] for l in range(n)])
    return out


def _sobol_tables(tmp_path):
    """The reference's Sobol' direction numbers, dumped by oracle/raydump.cpp `sobol` (test infrastructure)."""
    import subprocess
    out = str(tmp_path / "sobol.bin")
    r = subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", "sobol", out], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0 and os.path.exists(out), r.stdout[-2000:]
    return np.fromfile(out, np.uint32)


def _close(a, b, rtol, atol):
    return np.abs(a - b) <= atol + rtol * np.abs(b)


def _bsdf_pointwise(pkg, ctx, g, materials):
    S = pkg.scenes
    inp, ref = g["inp"], g["out"]
    sd = S.materials_scene(materials=materials)
    sc = sd.upload(ctx)
    try:
        for shape in range(len(materials)):
            rows = inp[:, 0] == shape
            bsdf_id = shape            # materials_scene adds one BSDF per shape in order
            out = sc.eval_bsdf(bsdf_id, inp[rows, 1:4], inp[rows, 4:7], inp[rows, 7:9], s3=inp[rows, 9])
            r = ref[rows]
            name = materials[shape]
            rtol = 2e-3 if name.get("distribution") == "beckmann" else 5e-4      # reference uses math::fastexp / erf fits
            ok_eval = _close(out[:, 0:4], r[:, 0:4], rtol, 1e-6).all(axis=1)
            assert ok_eval.mean() >= 0.995, (name, "eval/pdf", ok_eval.mean())
            # sampling: same lobe, same direction, same weight and density
            same_type = out[:, 12] == r[:, 12]
            assert same_type.mean() >= 0.995, (name, "sampledType", same_type.mean())
            m = same_type & (r[:, 4:7].sum(axis=1) > 0)
            ok_s = (_close(out[m, 4:7], r[m, 4:7], rtol * 2, 1e-5).all(axis=1) & _close(out[m, 7:10], r[m, 7:10], 0, 2e-4).all(axis=1) &
                    _close(out[m, 10], r[m, 10], rtol * 2, 1e-5) & _close(out[m, 11], r[m, 11], 1e-6, 0))
            assert ok_s.mean() >= 0.99, (name, "sample", ok_s.mean())
            zero = r[:, 4:7].sum(axis=1) == 0
            if zero.any():
                assert (out[zero, 4:7].sum(axis=1) == 0).mean() >= 0.99, (name, "failed samples")
    finally:
        sc.close()


def test_bsdf_pointwise_matches_reference_plugins(pkg, ctx, golden):
    """mtsb_eval_bsdf vs BSDF::eval / pdf / sample of diffuse, dielectric, plastic (linear, nonlinear), roughconductor
    (GGX visible, Beckmann visible, anisotropic GGX sampleAll), twosided, conductor and roughdielectric (GGX visible,
    anisotropic Beckmann sampleAll) — tests/golden/bsdf.npz."""
    _bsdf_pointwise(pkg, ctx, golden("bsdf.npz"), pkg.scenes.MATERIALS)


def test_roughplastic_pointwise_matches_reference_plugin(pkg, ctx, golden):
    """`roughplastic` (src/bsdfs/roughplastic.cpp:330-495): Beckmann / GGX, nonlinear, sampleVisible off, two-sided; the
    rough-transmittance slices are rebuilt by hostmath.rough_transmittance_slices from the reference's data/microfacet tables
    (rtrans.h:277-370) — tests/golden/bsdf_roughplastic.npz."""
    pkg.scenes.MICROFACET_DIR = os.path.join(O.REF_DIR, "data", "microfacet")
    _bsdf_pointwise(pkg, ctx, golden("bsdf_roughplastic.npz"), pkg.scenes.ROUGH_PLASTICS)


@pytest.mark.parametrize("which", ["emitter", "emitter_analytic", "emitter_envmap"])
def test_emitter_sampling_matches_reference(pkg, ctx, golden, which):
    """mtsb_eval_emitter vs Scene::sampleEmitterDirect (no visibility) + pdfEmitterDirect over a point light, a constant
    environment and a mesh area light — tests/golden/emitter.npz; and over area lights on analytic shapes: a sphere seen
    from outside (cone sampling), a sphere seen from inside, a disk (sphere.cpp:286-385, disk.cpp:247-259) —
    emitter_analytic.npz."""
    S = pkg.scenes
    g = golden(which + ".npz")
    inp, ref = g["inp"], g["out"]
    # emitter_envmap.npz: EnvironmentMap::sampleDirect / pdfDirect (envmap.cpp:525-606) on a rotated, scaled map beside a mesh light
    sd = S.materials_scene() if which == "emitter" else (S.analytic_lights_scene() if which == "emitter_analytic" else S.envmap_scene(pyramid=_pyramid(golden)))
    sc = sd.upload(ctx)
    try:
        out = sc.eval_emitter(inp[:, 0:3], inp[:, 3:6], inp[:, 6:8])
        same_zero = (out[:, 0:3].sum(axis=1) == 0) == (ref[:, 0:3].sum(axis=1) == 0)
        assert same_zero.mean() >= 0.995
        m = (ref[:, 0:3].sum(axis=1) > 0) & same_zero
        ok = (_close(out[m, 0:4], ref[m, 0:4], 1e-3, 1e-6).all(axis=1) & _close(out[m, 4:7], ref[m, 4:7], 0, 1e-4).all(axis=1) &
              (out[m, 8] == ref[m, 8]) & _close(out[m, 9], ref[m, 9], 1e-3, 1e-6))
        # dRec.dist — for the constant environment the distance to a bounding sphere built from Scene::getAABB() as it is
        # when the emitter's shape is created: geometry + the sensor's AABB (src/emitters/constant.cpp:69-80,
        # scene.cpp:394-404), which the scene description passes on with mtsb_scene_expand_bounds
        ok &= _close(out[m, 7], ref[m, 7], 1e-4, 1e-5)
        assert ok.mean() >= 0.995, ok.mean()
    finally:
        sc.close()


TAIL_ALL, TAIL_OFF = 1 << 30, 0     # every path through the late-bounce kernel / wavefront stages only (csrc/render_tail.cuh)


def _li_image(pkg, ctx, sd, tail=-1, positions=False):
    sc = sd.upload(ctx)
    w, h, spp = sd.sensor["width"], sd.sensor["height"], sd.sampler["spp"]
    rd = sd.render(sc)
    rd.set_tail(tail)
    ys, xs, ss = np.meshgrid(np.arange(h), np.arange(w), np.arange(spp), indexing="ij")
    samples = np.stack([xs.ravel(), ys.ravel(), ss.ravel(), np.zeros(xs.size)], axis=1).astype(np.uint32)
    if positions:
        L, pos = rd.render_samples(samples, spp, positions=True)
        L = np.concatenate([L, pos], axis=1).reshape(h, w, spp, 6)
    else:
        L = rd.render_samples(samples, spp).reshape(h, w, spp, 4)
    # the same samples through the film (ImageBlock layout, un-normalised sums)
    rd.render_pass(0, spp, spp)
    film = rd.film_to_host()
    stats = rd.stats()
    rd.close()
    sc.close()
    return L, film, stats


@pytest.mark.parametrize("tail", [TAIL_OFF, TAIL_ALL], ids=["wavefront", "tailkernel"])
@pytest.mark.parametrize("name", ["cornell", "materials", "materials_deep", "direct", "thinlens", "strict",
                                  "direct_4_2", "direct_1_3", "direct_3_1", "direct_2_0", "analytic", "roughplastic", "analytic_lights",
                                  "envmap", "envmap_direct", "textured", "textured_direct", "textured_plastic", "textured_thinlens",
                                  "ldctr_cornell", "ldctr_materials", "ldctr_direct"])
def test_li_matches_reference_per_sample(pkg, ctx, golden, name, tail):
    """Per-pixel mean of Li over the pixel's samples vs the reference's own `path` plugin rendering the same scene with
    the same random numbers (ctr sampler, box filter, classic block renderer) — once with every bounce run by the wavefront
    stages and once with every path carried by the late-bounce kernel."""
    S = pkg.scenes
    g = golden("li_%s.npz" % name)
    ref = g["image"]
    spp, seed = int(g["spp"]), int(g["seed"])
    if name == "cornell":
        sd = S.cornell_box(32, 32, spp=spp, max_depth=5, sampler="ctr", seed=seed)
    elif name.startswith("ldctr_"):
        # the ldsampler's numbers (csrc/shading.cuh ld_next1D / ld_next2D) against the reference integrators fed the same numbers
        # by the counter sampler's low-discrepancy twin (oracle/ctrsampler.cpp): the 1D and 2D counters advance separately, so
        # this pins which draws are 1D and which are 2D; parameters of make_golden_render.ldctr_cases
        if name == "ldctr_cornell":
            sd = S.cornell_box(32, 32, spp=spp, max_depth=5, sampler="ldctr", seed=seed)
        else:
            sd = S.materials_scene(48, 32, spp=spp, max_depth=6, sampler="ldctr", seed=seed)
            if name == "ldctr_direct":
                sd.sampler["dimension"] = 2
                sd.set_integrator("direct")
    elif name == "materials":
        sd = S.materials_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed)
    elif name == "direct":
        # MIDirectIntegrator::Li (src/integrators/direct/direct.cpp:146-309) with emitterSamples = bsdfSamples = 1
        sd = S.materials_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed)
        sd.set_integrator("direct")
    elif name.startswith("direct_"):
        # emitterSamples / bsdfSamples other than 1/1 (direct.cpp:95-107, 195-306): sample arrays, fraction-weighted power
        # heuristic; parameters of make_golden_render.direct_fan_case
        n_em, n_bsdf = (int(x) for x in name.split("_")[1:])
        sd = S.materials_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed)
        sd.set_integrator("direct", emitter_samples=n_em, bsdf_samples=n_bsdf)
    elif name == "analytic":
        # analytic sphere / disk shapes (src/shapes/sphere.cpp, disk.cpp) through the path tracer
        sd = S.analytic_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed)
    elif name in ("envmap", "envmap_direct"):
        # `envmap` emitter: importance sampling, bilinear lookups for scattered rays, EWA-filtered lookups (ray differentials scaled
        # by 1 / sqrt(spp)) where the camera sees the map directly — through `path` and through `direct`
        sd = S.envmap_scene(48, 32, spp=spp, max_depth=5, sampler="ctr", seed=seed, pyramid=_pyramid(golden),
                            kind="direct" if name == "envmap_direct" else "path")
    elif name == "textured_thinlens":
        # ray differentials of the thinlens sensor (thinlens.cpp:324-372) feeding the texture lookups of the first hit;
        # parameters of make_golden_render.textured_thinlens
        sd = S.textured_scene(48, 32, spp=spp, max_depth=5, sampler="ctr", seed=seed, pyramids=_tex_pyramids(golden))
        s = sd.sensor
        sd.set_sensor(s["to_world"], s["fov"], s["width"], s["height"], rfilter="box", spp=spp, seed=seed, sampler="ctr",
                      aperture_radius=0.08, focus_distance=4.6)
    elif name in ("textured", "textured_direct", "textured_plastic"):
        # `bitmap` textures on diffuse reflectances: uv partials from the sensor ray's differentials at the first hit
        # (intersection.cpp:5-76) and EWA / trilinear lookups there, full-resolution lookups at later hits; repeat / clamp / mirror
        sd = S.textured_scene(48, 32, spp=spp, max_depth=5, sampler="ctr", seed=seed, pyramids=_tex_pyramids(golden),
                              kind="direct" if name == "textured_direct" else "path", plastic=name == "textured_plastic")
    elif name == "analytic_lights":
        # area lights on analytic spheres (cone / uniform-sphere sampling, sphere.cpp:286-385) and a disk (disk.cpp:247-259)
        sd = S.analytic_lights_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed)
    elif name == "roughplastic":
        S.MICROFACET_DIR = os.path.join(O.REF_DIR, "data", "microfacet")
        sd = S.materials_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed, materials=S.ROUGH_PLASTICS)
    elif name == "strict":
        # strictNormals + hideEmitters (path.cpp:136-165,196-199,218-222); parameters of make_golden_render.strict_case
        sd = S.materials_scene(48, 32, spp=spp, max_depth=5, sampler="ctr", seed=seed)
        sd.set_integrator("path", max_depth=5, rr_depth=3, strict_normals=True, hide_emitters=True)
    elif name == "thinlens":
        # ThinLens::sampleRay (src/sensors/thinlens.cpp:293-322); the parameters are those of make_golden_render.thinlens_case
        sd = S.materials_scene(48, 32, spp=spp, max_depth=4, sampler="ctr", seed=seed)
        s = sd.sensor
        sd.set_sensor(s["to_world"], s["fov"], s["width"], s["height"], rfilter="box", spp=spp, seed=seed, sampler="ctr",
                      aperture_radius=0.12, focus_distance=5.2)
    else:
        sd = S.materials_scene(24, 16, spp=spp, max_depth=-1, sampler="ctr", seed=seed)
    L, film, stats = _li_image(pkg, ctx, sd, tail=tail)
    img = L[..., 0:3].mean(axis=2)
    ok = _close(img, ref, 2e-3, 2e-4).all(axis=2)
    # a path whose hit flips across a triangle edge (rounding) takes a different course: bounded fraction of pixels
    assert ok.mean() >= 0.97, "only %.4f of the pixels agree" % ok.mean()
    assert abs(img.mean() - ref.mean()) <= 0.01 * ref.mean()
    # film == sum of filter-weighted samples: box filter -> weight table[0]^2 on the sample's own pixel
    b, ch = sd.film().border, 5
    wgt = film[b:film.shape[0] - b, b:film.shape[1] - b, 4]
    # (a sample within 1e-5 of a pixel boundary also reaches the neighbour, exactly as in ImageBlock::put with the box
    # filter's radius of 0.5 + 1e-5, src/rfilters/box.cpp:33-40 — those few pixels are left out of the comparison)
    nominal = np.isclose(wgt, np.median(wgt), rtol=1e-6)
    assert nominal.mean() >= 0.995 and abs(np.median(wgt) / spp - 1) < 1e-3
    dev = film[b:film.shape[0] - b, b:film.shape[1] - b, 0:3] / wgt[..., None]
    assert np.allclose(dev[nominal], img[nominal], rtol=1e-4, atol=1e-5)
    assert stats["paths"] == img.shape[0] * img.shape[1] * spp * 2      # render_samples + render_pass
    assert stats["closest_rays"] >= stats["paths"] and stats["depth_sum"] >= stats["paths"]


@pytest.mark.parametrize("name", ["fog", "fog_short", "fog_haze", "fogmis", "fogmis_short", "fogmis_haze", "fog_ldctr", "fogmis_ldctr", "fog_sphere", "fogmis_sphere", "fog_envmap", "fogmis_envmap",
                                  "fog_strict", "fogmis_strict", "fog_plastic_thinlens", "fogmis_plastic_thinlens"])
def test_volpath_simple_matches_reference_per_sample(pkg, ctx, golden, name):
    """SimpleVolumetricPathTracer::Li (src/integrators/path/volpath_simple.cpp:88-300) with homogeneous media
    (src/medium/homogeneous.cpp), isotropic / HG phase functions and Scene::evalTransmittance through index-matched
    boundaries (scene.cpp:650-713), against the reference's own plugins on the same random numbers; parameters of
    make_golden_render.fog_cases."""
    S = pkg.scenes
    g = golden("li_%s.npz" % name)
    ref = g["image"]
    spp, seed = int(g["spp"]), int(g["seed"])
    # fogmis*: VolumetricPathTracer::Li (volpath.cpp:84-425), the variant with multiple importance sampling
    kind = "volpath" if name.startswith("fogmis") else "volpath_simple"
    if name.endswith("_sphere"):
        # an analytic sphere light in the haze: volshade_kernel<*, PLAIN = false> (the variants with the analytic-light, texture and
        # environment-map code compiled in), which the other fog scenes never launch
        sd = S.fog_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed, haze=True, sphere_light=True, kind=kind)
    elif name.endswith("_plastic_thinlens"):
        # a plastic ball: volshade_kernel<*, PLAIN, LEAN = false> under the independent sampler; thinlens sensor inside the haze
        sd = S.fog_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed, haze=True, kind=kind, aperture_radius=0.1,
                         ball_bsdf=dict(type="plastic", diffuse_reflectance=(0.6, 0.3, 0.2), nonlinear=True))
    elif name.endswith("_strict"):
        # strictNormals + hideEmitters (volpath_simple.cpp:176-197,223,266); parameters of make_golden_render.fog_cases
        sd = S.fog_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed, haze=True, kind=kind)
        sd.set_integrator(kind, max_depth=6, rr_depth=3, strict_normals=True, hide_emitters=True)
    elif name.endswith("_envmap"):
        # the environment map seen and sampled through the media (importance sampling, EWA lookups where the camera sees it)
        sd = S.fog_scene(48, 32, spp=spp, max_depth=6, sampler="ctr", seed=seed, haze=True, envmap=True, envmap_pyramid=_pyramid(golden), kind=kind)
    elif name.endswith("_ldctr"):
        # the ldsampler's numbers through the reference's volumetric integrators (the counter sampler's low-discrepancy twin):
        # distances and lobe choices are 1D draws, phase functions and emitter samples 2D draws, counted separately
        sd = S.fog_scene(48, 32, spp=spp, max_depth=6, sampler="ldctr", seed=seed, haze=True, kind=kind)
    else:
        sd = S.fog_scene(48, 32, spp=spp, max_depth={"": 6, "_short": 3, "_haze": -1}[name.replace("fogmis", "").replace("fog", "")], sampler="ctr", seed=seed,
                         haze=name.endswith("_haze"), kind=kind)
    L, film, stats = _li_image(pkg, ctx, sd)
    img = L[..., 0:3].mean(axis=2)
    ok = _close(img, ref, 2e-3, 2e-4).all(axis=2)
    print(name, "pixels within tolerance: %.4f" % ok.mean(), "mean", img.mean(), "ref", ref.mean())
    assert ok.mean() >= 0.97, "only %.4f of the pixels agree" % ok.mean()
    assert abs(img.mean() - ref.mean()) <= 0.01 * ref.mean()
    # opacity (records.inl:117-145): 1 on opaque first hits, inside [0, 1] behind index-matched boundaries / in the haze
    a = L[..., 3]
    assert np.isfinite(a).all() and a.min() >= 0.0 and a.max() <= 1.0      # a negative channel would be rejected by ImageBlock::put
    assert (a == 1.0).mean() > 0.3      # an opaque surface behind a fog ball makes the pixel opaque (T = 0 at the occluder)
    b = sd.film().border
    wgt = film[b:film.shape[0] - b, b:film.shape[1] - b, 4]
    nominal = np.isclose(wgt, np.median(wgt), rtol=1e-6)
    dev = film[b:film.shape[0] - b, b:film.shape[1] - b, 0:3] / wgt[..., None]
    assert np.allclose(dev[nominal], img[nominal], rtol=1e-4, atol=1e-5)
    assert stats["paths"] == img.shape[0] * img.shape[1] * spp * 2


@pytest.mark.parametrize("name", ["sobol_materials", "sobol_cornell", "sobol_fog", "sobol_fogmis"])
def test_sobol_sampler_matches_reference_per_sample(pkg, ctx, golden, tmp_path, name):
    """The `sobol` sampler (src/samplers/sobol.cpp:172-251 over sobolseq.h): pixel-wise enumeration of the sequence (look_up on the
    film size rounded up to a power of two), the first two dimensions as the position inside the pixel, one dimension per later
    draw — against images the reference's own sobol plugin rendered through `path` (make_golden_render.sobol_cases)."""
    S = pkg.scenes
    S.SOBOL_TABLES = _sobol_tables(tmp_path)
    g = golden("li_%s.npz" % name)
    ref = g["image"]
    spp = int(g["spp"])
    if name in ("sobol_fog", "sobol_fogmis"):       # volpath_simple / volpath over the haze variant of the fog scene
        sd = S.fog_scene(48, 32, spp=spp, max_depth=6, sampler="sobol", seed=0, haze=True, kind="volpath" if name == "sobol_fogmis" else "volpath_simple")
    else:
        sd = (S.materials_scene(48, 32, spp=spp, max_depth=6, sampler="sobol", seed=0) if name == "sobol_materials"
              else S.cornell_box(40, 24, spp=spp, max_depth=5, sampler="sobol", seed=0))
    for tail in (TAIL_OFF, TAIL_ALL) if "fog" not in name else (TAIL_OFF,):      # the volumetric integrators have no tail kernel
        L, film, stats = _li_image(pkg, ctx, sd, tail=tail)
        # Sobol' points are dyadic: one pixel in 64 owns a sample exactly on its corner, which the box filter (radius 0.5 + 1e-5,
        # src/rfilters/box.cpp:33-40) also splats into the three pixels that share the corner.  The reference image is a developed
        # film, so the comparison is film against film (ImageBlock::put on both sides) ...
        b = sd.film().border
        wgt = film[b:film.shape[0] - b, b:film.shape[1] - b, 4]
        dev = film[b:film.shape[0] - b, b:film.shape[1] - b, 0:3] / wgt[..., None]
        ok = _close(dev, ref, 2e-3, 2e-4).all(axis=2)
        print(name, "tail", tail, "pixels within tolerance: %.4f" % ok.mean(), "pixels with a neighbour's corner sample: %.4f" % (~np.isclose(wgt, np.median(wgt), rtol=1e-6)).mean())
        assert ok.mean() >= 0.97, "only %.4f of the pixels agree" % ok.mean()
        assert abs(dev.mean() - ref.mean()) <= 0.01 * ref.mean()
        # ... and the per-sample radiances average to the film wherever a pixel holds only its own samples
        img = L[..., 0:3].mean(axis=2)
        nominal = np.isclose(wgt, np.median(wgt), rtol=1e-6)
        assert nominal.mean() > 0.9 and np.allclose(dev[nominal], img[nominal], rtol=1e-4, atol=1e-5)


def _imageblock_put(shape, border, radius, table, pos, values):
    """ImageBlock::put(pos, value) restated in numpy (include/mitsuba/render/imageblock.h:161-221): offset 0, the filter's
    32-entry table through evalDiscretized (rfilter.h:76-77), float64 accumulation."""
    out = np.zeros(shape, np.float64)
    size_y, size_x = shape[0], shape[1]
    scale = np.float32(31.0) / np.float32(radius)
    px = pos[:, 0].astype(np.float32) - np.float32(0.5) - np.float32(-border)
    py = pos[:, 1].astype(np.float32) - np.float32(0.5) - np.float32(-border)
    r = np.float32(radius)
    x0 = np.maximum(np.ceil(px - r).astype(np.int64), 0); x1 = np.minimum(np.floor(px + r).astype(np.int64), size_x - 1)
    y0 = np.maximum(np.ceil(py - r).astype(np.int64), 0); y1 = np.minimum(np.floor(py + r).astype(np.int64), size_y - 1)
    span = int(np.ceil(2 * radius)) + 1
    for dy in range(span):
        y = y0 + dy
        wy = table[np.minimum((np.abs((y.astype(np.float32) - py)) * scale).astype(np.int64), 31)]
        for dx in range(span):
            x = x0 + dx
            ok = (y <= y1) & (x <= x1)
            wx = table[np.minimum((np.abs((x.astype(np.float32) - px)) * scale).astype(np.int64), 31)]
            w = (wx * wy).astype(np.float32)
            np.add.at(out, (y[ok], x[ok]), (w[ok, None] * values[ok]).astype(np.float64))
    return out


@pytest.mark.parametrize("rfilter", ["gaussian", "box"])
def test_film_splat_matches_imageblock_put_per_pixel(pkg, ctx, rfilter):
    """The device film equals, pixel by pixel and channel by channel, ImageBlock::put applied on the host to the very
    samples the device rendered (position px + u and radiance of every sample come back through
    mtsb_render_samples_ex).  The Gaussian filter (sigma 0.5, radius 2, border 2) is the one configs 3-5 use; its
    warp-tiled shared-memory accumulation (render_shade.cuh: splat_kernel) is what this pins."""
    S = pkg.scenes
    sd = S.cornell_box(40, 28, spp=3, max_depth=3, sampler="ctr", seed=5, rfilter=rfilter)
    L, film, _ = _li_image(pkg, ctx, sd, positions=True)
    fd = sd.film()
    r, table, border = S.filter_table(rfilter)
    assert border == fd.border
    h, w, spp = L.shape[0:3]
    assert film.shape == (h + 2 * border, w + 2 * border, 5)
    flat = L.reshape(-1, 6)
    values = np.concatenate([flat[:, 0:4], np.ones((flat.shape[0], 1), np.float32)], axis=1).astype(np.float32)
    ref = _imageblock_put(film.shape, border, r, np.asarray(table, np.float32), flat[:, 4:6], values)
    # float32 atomics in a different order than the host's float64 sums: 1e-5 relative to the pixel's own magnitude
    tol = 1e-5 * np.maximum(np.abs(ref), 1.0) + 1e-6
    bad = np.abs(film - ref) > tol
    assert not bad.any(), "%d of %d film floats differ, worst %.3g" % (bad.sum(), bad.size, np.abs(film - ref).max())
    assert abs(film[..., 4].sum() - ref[..., 4].sum()) <= 1e-5 * ref[..., 4].sum()


def test_headless_preview_stack_matches_the_preview_arithmetic(pkg, ctx, tmp_path):
    """mtsb_preview_stack against the stacking arithmetic of the im-mts preview (src/im-mts/preview_opengl2.cpp:188-270,
    additive blend with constant colour 1 / max(sum spp, 1) over alpha 1) restated in numpy, on two worker films that hold
    the interleaved sample indices of one frame — and the frame lands in a PFM file, as a headless preview has to."""
    S, B = pkg.scenes, pkg.binding
    sd = S.cornell_box(48, 32, spp=6, max_depth=4, sampler="ctr", seed=9, rfilter="gaussian")
    sc = sd.upload(ctx)
    try:
        renders = [B.Render(sc, sd.camera(), sd.film(channels=4), sd.integrator_desc(), seed=9) for _ in range(2)]
        for k, rd in enumerate(renders):
            rd.render_pass(0, 3, 6, stride=2, phase=k)           # worker k: sample indices k, k + 2, k + 4
        frame = B.preview_stack(renders, [3.0, 3.0])
        films = [rd.film_to_host().astype(np.float64) for rd in renders]
        ref = (films[0][..., 0:3] + films[1][..., 0:3]) / 6.0
        assert frame.shape == films[0].shape[0:2] + (4,)
        assert np.allclose(frame[..., 0:3], ref, rtol=1e-6, atol=1e-7) and (frame[..., 3] == 1).all()
        # fewer than one sample per pixel on average: the scale is clamped at 1 (std::max(avgSamples, 1.f))
        half = B.preview_stack(renders[0:1], [0.5])
        assert np.allclose(half[..., 0:3], films[0][..., 0:3], rtol=1e-6, atol=1e-7)
        # the frame as a file
        path = str(tmp_path / "frame.pfm")
        with open(path, "wb") as f:
            f.write(b"PF\n%d %d\n-1.0\n" % (frame.shape[1], frame.shape[0]))
            f.write(np.ascontiguousarray(frame[::-1, :, 0:3], "<f4").tobytes())
        assert np.allclose(O.read_pfm(path), frame[..., 0:3])
        for rd in renders:
            rd.close()
    finally:
        sc.close()


def test_image_is_unbiased_against_live_reference(pkg, ctx, tmp_path):
    """relMSE of the GPU render against a high-spp render of the unmodified reference shrinks like 1/spp (unbiased
    estimator), and the converged means agree — materials scene, independent reference sampler."""
    if not O.have_reference():
        pytest.fail("oracle/_ref is missing")
    S = pkg.scenes
    w, h = 64, 48
    sd_ref = S.materials_scene(w, h, spp=1024, max_depth=6, sampler="independent")
    xml = sd_ref.to_xml(str(tmp_path), "mats_ref")
    pfm = str(tmp_path / "ref.pfm")
    O.ref_render(xml, pfm, threads=os.cpu_count() or 1, classic=True)
    ref = O.read_pfm(pfm).astype(np.float64)
    sc = sd_ref.upload(ctx)
    rel = {}
    try:
        for spp in (4, 16, 64, 256):
            rd = pkg.binding.Render(sc, sd_ref.camera(), sd_ref.film(), sd_ref.integrator_desc(), seed=1234 + spp)
            rd.render_pass(0, spp, spp)
            film = rd.film_to_host().astype(np.float64)
            b = sd_ref.film().border
            film = film[b:film.shape[0] - b, b:film.shape[1] - b]
            rd.close()
            img = film[..., 0:3] / film[..., 4:5]
            rel[spp] = float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))
        print("relMSE vs 1024-spp reference:", rel)
        assert rel[16] < rel[4] * 0.5 and rel[64] < rel[16] * 0.5 and rel[256] < rel[64] * 0.6
        # tolerance that shrinks with spp: relMSE(spp) <= 1.5 * relMSE(4) * 4 / spp + reference noise floor
        floor = rel[4] * 4 / 1024 * 2
        for spp in (16, 64, 256):
            assert rel[spp] <= 1.5 * rel[4] * 4 / spp + floor, (spp, rel)
        assert abs(img.mean() - ref.mean()) <= 0.02 * ref.mean()
    finally:
        sc.close()
