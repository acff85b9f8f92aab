"""Golden vectors for the shading half of the hot path, produced by the UNMODIFIED reference (oracle/_ref):

  bsdf.npz     BSDF::eval / pdf / sample of every BSDF in scenes.MATERIALS on random directions (oracle/raydump.cpp `bsdf`)
  emitter.npz  Scene::sampleEmitterDirect / pdfEmitterDirect on random reference points (`emitter`)
  li_*.npz     images rendered by the reference's own `path` plugin through `mitsuba -C` with the counter-based `ctr`
               sampler (oracle/ctrsampler.cpp) and a box filter: each pixel is the mean of MIPathTracer::Li over the
               pixel's samples, which the CUDA path must reproduce sample by sample.

Run here:  python tests/golden/make_golden_render.py
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as g  # noqa: E402
import oracle as O  # noqa: E402

pkg = g.load_package()
S = pkg.scenes
S.MICROFACET_DIR = os.path.join(O.REF_DIR, "data", "microfacet")
GOLD = os.path.join(ROOT, "tests", "golden")


def run_util(mode, xml, inp, ncols_out):
    with tempfile.TemporaryDirectory() as td:
        ip, op = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        np.ascontiguousarray(inp, np.float32).tofile(ip)
        r = subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", mode, xml, ip, op], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0 or not os.path.exists(op):
            raise RuntimeError(r.stdout[-3000:])
        return np.fromfile(op, np.float32).reshape(-1, ncols_out)


def rand_dirs(rng, n, upper_frac=0.8):
    v = rng.normal(size=(n, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    flip = rng.random(n) < upper_frac
    v[:, 2] = np.where(flip, np.abs(v[:, 2]), v[:, 2])
    return v.astype(np.float32)


def bsdf_case(td):
    sd = S.materials_scene()
    xml = sd.to_xml(td, "mats")
    rng = np.random.default_rng(5)
    n = 384
    rows = []
    for shape in range(len(S.MATERIALS)):
        inp = np.zeros((n, 12), np.float32)
        inp[:, 0] = shape
        inp[:, 1:4] = rand_dirs(rng, n)
        inp[:, 4:7] = rand_dirs(rng, n, 0.7)
        inp[:, 7:9] = rng.random((n, 2), dtype=np.float32)
        inp[:, 9] = rng.random(n, dtype=np.float32)     # the draw EUsesSampler plugins take inside sample()
        rows.append(inp)
    inp = np.concatenate(rows)
    out = run_util("bsdf", xml, inp, 16)
    np.savez_compressed(os.path.join(GOLD, "bsdf.npz"), inp=inp, out=out)
    print("bsdf:", inp.shape, "nonzero eval frac", (out[:, 0:3].sum(1) > 0).mean(), "nonzero sample frac", (out[:, 4:7].sum(1) > 0).mean())


def emitter_case(td):
    sd = S.materials_scene()
    xml = sd.to_xml(td, "mats_e")
    rng = np.random.default_rng(6)
    n = 1536
    inp = np.zeros((n, 8), np.float32)
    inp[:, 0:3] = rng.uniform([-2, 0, -1.5], [2, 1.5, 1.5], (n, 3))
    nrm = rand_dirs(rng, n, 0.5)
    nrm[: n // 4] = 0          # refN = 0: transmissive reference surface
    inp[:, 3:6] = nrm
    inp[:, 6:8] = rng.random((n, 2), dtype=np.float32)
    out = run_util("emitter", xml, inp, 12)
    np.savez_compressed(os.path.join(GOLD, "emitter.npz"), inp=inp, out=out)
    print("emitter:", inp.shape, "nonzero frac", (out[:, 0:3].sum(1) > 0).mean(), "measures", np.unique(out[:, 8]))


def li_case(td, name, sd):
    xml = sd.to_xml(td, name)
    pfm = os.path.join(td, name + ".pfm")
    O.ref_render(xml, pfm, threads=4, classic=True)
    img = O.read_pfm(pfm)
    np.savez_compressed(os.path.join(GOLD, "li_%s.npz" % name), image=img, spp=sd.sampler["spp"], seed=sd.sampler["seed"])
    print(name, img.shape, "mean", img.mean(axis=(0, 1)))


def direct_case(td):
    """MIDirectIntegrator (src/integrators/direct/direct.cpp) with its default emitterSamples = bsdfSamples = 1."""
    sd = S.materials_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=13)
    sd.set_integrator("direct")
    li_case(td, "direct", sd)


def direct_fan_case(td, n_em, n_bsdf):
    """MIDirectIntegrator with emitterSamples / bsdfSamples other than 1: the sample ARRAYS of the sampler come into play
    (direct.cpp:140-145, 200-204, 250-254), which the counter sampler defines as a second Philox stream."""
    sd = S.materials_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=29)
    sd.set_integrator("direct", emitter_samples=n_em, bsdf_samples=n_bsdf)
    li_case(td, "direct_%d_%d" % (n_em, n_bsdf), sd)


def strict_case(td):
    """MonteCarloIntegrator's strictNormals and hideEmitters switches (src/librender/integrator.cpp:191-226,
    src/integrators/path/path.cpp:136-165,196-199,218-222) on meshes with interpolated normals."""
    sd = S.materials_scene(48, 32, spp=4, max_depth=5, sampler="ctr", seed=23)
    sd.set_integrator("path", max_depth=5, rr_depth=3, strict_normals=True, hide_emitters=True)
    li_case(td, "strict", sd)


def thinlens_case(td):
    """ThinLens::sampleRay (src/sensors/thinlens.cpp:293-322): the aperture sample is drawn right after the pixel sample."""
    sd = S.materials_scene(48, 32, spp=4, max_depth=4, sampler="ctr", seed=17)
    s = sd.sensor
    sd.set_sensor(s["to_world"], s["fov"], s["width"], s["height"], rfilter="box", spp=4, seed=17, sampler="ctr",
                  aperture_radius=0.12, focus_distance=5.2)
    li_case(td, "thinlens", sd)


def roughplastic_cases(td):
    """`roughplastic` (src/bsdfs/roughplastic.cpp + the RoughTransmittance tables of src/bsdfs/rtrans.h): pointwise
    eval / pdf / sample of scenes.ROUGH_PLASTICS and a per-sample render of the same spheres."""
    sd = S.materials_scene(materials=S.ROUGH_PLASTICS)
    xml = sd.to_xml(td, "rp")
    rng = np.random.default_rng(8)
    n = 512
    rows = []
    for shape in range(len(S.ROUGH_PLASTICS)):
        inp = np.zeros((n, 12), np.float32)
        inp[:, 0] = shape
        inp[:, 1:4] = rand_dirs(rng, n)
        inp[:, 4:7] = rand_dirs(rng, n, 0.9)
        inp[:, 7:9] = rng.random((n, 2), dtype=np.float32)
        inp[:, 9] = rng.random(n, dtype=np.float32)
        rows.append(inp)
    inp = np.concatenate(rows)
    out = run_util("bsdf", xml, inp, 16)
    np.savez_compressed(os.path.join(GOLD, "bsdf_roughplastic.npz"), inp=inp, out=out)
    print("bsdf_roughplastic:", inp.shape, "nonzero eval frac", (out[:, 0:3].sum(1) > 0).mean(), "nonzero sample frac", (out[:, 4:7].sum(1) > 0).mean())
    li_case(td, "roughplastic", S.materials_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=47, materials=S.ROUGH_PLASTICS))


def analytic_lights_cases(td):
    """Area lights on analytic spheres / disks (src/shapes/sphere.cpp:286-385, disk.cpp:247-259 under AreaLight): pointwise
    Scene::sampleEmitterDirect / pdfEmitterDirect from points outside and inside the spheres, and a per-sample render."""
    sd = S.analytic_lights_scene()
    xml = sd.to_xml(td, "alights")
    rng = np.random.default_rng(16)
    n = 2048
    inp = np.zeros((n, 8), np.float32)
    inp[:, 0:3] = rng.uniform([-2.5, 0, -2], [2.5, 3.0, 2], (n, 3))
    nrm = rand_dirs(rng, n, 0.5)
    nrm[: n // 4] = 0
    inp[:, 3:6] = nrm
    inp[:, 6:8] = rng.random((n, 2), dtype=np.float32)
    out = run_util("emitter", xml, inp, 12)
    np.savez_compressed(os.path.join(GOLD, "emitter_analytic.npz"), inp=inp, out=out)
    print("emitter_analytic:", inp.shape, "nonzero frac", (out[:, 0:3].sum(1) > 0).mean())
    li_case(td, "analytic_lights", S.analytic_lights_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=59))


def envmap_cases(td):
    """`envmap` (src/emitters/envmap.cpp): the MIP pyramid the reference builds from scenes.sky_image() (oracle/raydump.cpp
    `mipmap`), pointwise Scene::sampleEmitterDirect / pdfEmitterDirect, and per-sample renders through `path` and `direct` (the
    camera sees the map directly: the EWA-filtered lookups of sensor rays, mipmap.h:630-716)."""
    img = S.sky_image()
    S.write_pfm(os.path.join(td, "sky.pfm"), img)
    mp = os.path.join(td, "mip.bin")
    r = subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", "mipmap", os.path.join(td, "sky.pfm"), mp], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0 or not os.path.exists(mp):
        raise RuntimeError(r.stdout[-3000:])
    a = np.fromfile(mp, np.float32)
    levels, off, pyr = int(a[0]), 1, {}
    for l in range(levels):
        w, h = int(a[off]), int(a[off + 1]); off += 2
        pyr["level%d" % l] = a[off:off + w * h * 3].reshape(h, w, 3).astype(np.float16)      # the reference stores half precision
        assert (pyr["level%d" % l].astype(np.float32) == a[off:off + w * h * 3].reshape(h, w, 3)).all()
        off += w * h * 3
    np.savez_compressed(os.path.join(GOLD, "envmap_pyramid.npz"), **pyr)
    pyramid = [pyr["level%d" % l].astype(np.float32) for l in range(levels)]
    sd = S.envmap_scene(pyramid=pyramid)
    xml = sd.to_xml(td, "env")
    rng = np.random.default_rng(21)
    n = 2048
    inp = np.zeros((n, 8), np.float32)
    inp[:, 0:3] = rng.uniform([-2, 0, -1.5], [2, 1.5, 1.5], (n, 3))
    nrm = rand_dirs(rng, n, 0.5)
    nrm[: n // 4] = 0
    inp[:, 3:6] = nrm
    inp[:, 6:8] = rng.random((n, 2), dtype=np.float32)
    out = run_util("emitter", xml, inp, 12)
    np.savez_compressed(os.path.join(GOLD, "emitter_envmap.npz"), inp=inp, out=out)
    print("emitter_envmap:", inp.shape, "nonzero frac", (out[:, 0:3].sum(1) > 0).mean(), "emitters", np.unique(out[:, 8]))
    li_case(td, "envmap", S.envmap_scene(48, 32, spp=4, max_depth=5, sampler="ctr", seed=67, pyramid=pyramid))
    li_case(td, "envmap_direct", S.envmap_scene(48, 32, spp=4, max_depth=5, sampler="ctr", seed=71, pyramid=pyramid, kind="direct"))


def texture_cases(td):
    """`bitmap` textures (src/textures/bitmap.cpp): the MIP pyramids the reference builds from scenes.texture_image() under the three
    wrap modes of scenes.textured_scene (oracle/raydump.cpp `mipmap3`), and per-sample renders through `path` and `direct`."""
    img = S.texture_image()
    S.write_pfm(os.path.join(td, "tex.pfm"), img)
    B = pkg.binding
    pyramids, store = [], {}
    for k, wrap in enumerate(("repeat", "clamp", "mirror")):
        mp = os.path.join(td, "mip3_%d.bin" % k)
        w = str(B.WRAP_MODES[wrap])
        r = subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", "mipmap3", os.path.join(td, "tex.pfm"), mp, w, w], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0 or not os.path.exists(mp):
            raise RuntimeError(r.stdout[-3000:])
        a = np.fromfile(mp, np.float32)
        levels, off, pyr = int(a[0]), 1, []
        for l in range(levels):
            w_, h_ = int(a[off]), int(a[off + 1]); off += 2
            pyr.append(a[off:off + w_ * h_ * 3].reshape(h_, w_, 3).copy()); off += w_ * h_ * 3
            store["%s_level%d" % (wrap, l)] = pyr[-1]
        pyramids.append(pyr)
    np.savez_compressed(os.path.join(GOLD, "texture_pyramids.npz"), **store)
    li_case(td, "textured", S.textured_scene(48, 32, spp=4, max_depth=5, sampler="ctr", seed=79, pyramids=pyramids))
    li_case(td, "textured_direct", S.textured_scene(48, 32, spp=4, max_depth=5, sampler="ctr", seed=83, pyramids=pyramids, kind="direct"))
    li_case(td, "textured_plastic", S.textured_scene(48, 32, spp=4, max_depth=5, sampler="ctr", seed=89, pyramids=pyramids, plastic=True))
    li_case(td, "textured_thinlens", textured_thinlens(pyramids))


def textured_thinlens(pyramids, spp=4, seed=113):
    """The textured scene through a `thinlens` sensor: ThinLensCamera::sampleRayDifferential (thinlens.cpp:324-372) aims the two
    offset rays through the SAME aperture point, which is what the uv partials of the first hit are computed from."""
    sd = S.textured_scene(48, 32, spp=spp, max_depth=5, sampler="ctr", seed=seed, pyramids=pyramids)
    s = sd.sensor
    sd.set_sensor(s["to_world"], s["fov"], s["width"], s["height"], rfilter="box", spp=spp, seed=seed, sampler="ctr",
                  aperture_radius=0.08, focus_distance=4.6)
    return sd


def ldctr_cases(td):
    """The low-discrepancy twin of the counter sampler (oracle/ctrsampler.cpp, ldDimension > 0): the reference's `path` / `direct`
    consume the CUDA path's ldsampler numbers draw for draw — pins WHICH draws are 1D and which are 2D, which the shared
    independent stream cannot tell apart."""
    sd = S.cornell_box(32, 32, spp=8, max_depth=5, sampler="ldctr", seed=97)
    li_case(td, "ldctr_cornell", sd)
    sd = S.materials_scene(48, 32, spp=4, max_depth=6, sampler="ldctr", seed=101)
    li_case(td, "ldctr_materials", sd)
    sd = S.materials_scene(48, 32, spp=4, max_depth=6, sampler="ldctr", seed=103)
    sd.sampler["dimension"] = 2
    sd.set_integrator("direct")
    li_case(td, "ldctr_direct", sd)
    # the volumetric integrators draw 1D numbers for distances and lobes, 2D numbers for phase functions and emitters
    li_case(td, "fog_ldctr", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ldctr", seed=107, haze=True))
    li_case(td, "fogmis_ldctr", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ldctr", seed=109, haze=True, kind="volpath"))


def sobol_cases(td):
    """The `sobol` sampler (src/samplers/sobol.cpp): the reference's own plugin drives `path` on the materials scene (48 x 32 -> the
    enumeration works on a 64 x 64 grid) and on the Cornell box at a non-power-of-two size; deterministic, no shared sampler needed."""
    li_case(td, "sobol_materials", S.materials_scene(48, 32, spp=4, max_depth=6, sampler="sobol", seed=0))
    li_case(td, "sobol_cornell", S.cornell_box(40, 24, spp=8, max_depth=5, sampler="sobol", seed=0))
    # through the volumetric integrators: distance / lobe draws interleave 1D and 2D requests, which moves the places where the
    # reserved-dimension quirk of next2D bites
    li_case(td, "sobol_fog", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="sobol", seed=0, haze=True))
    li_case(td, "sobol_fogmis", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="sobol", seed=0, haze=True, kind="volpath"))


def all_case(td):
    """Scene::rayIntersectAll x2 (src/librender/scene.cpp:716-792; oracle/raydump.cpp `all`): the kd-tree shapes plus the special
    shapes of Scene::initializeBidirectional — here the aperture disk of the thinlens sensor and the bounding sphere of the
    constant environment.  Rays: chords through the scene (they end on the environment sphere when they miss the geometry),
    rays aimed at the aperture, and short segments that stop before anything."""
    sd = S.materials_scene(48, 32, spp=1, max_depth=4, sampler="ctr", seed=1)
    s = sd.sensor
    sd.set_sensor(s["to_world"], s["fov"], s["width"], s["height"], rfilter="box", spp=1, seed=1, sampler="ctr",
                  aperture_radius=0.12, focus_distance=5.2)
    xml = sd.to_xml(td, "all")
    rng = np.random.default_rng(12)
    n = 6000
    rays = np.zeros((n, 8), np.float32)
    a = S.chord_rays(n, center=(0, 0.5, 0), radius=5.0, seed=77, mint=S.EPSILON)
    rays[:, 0:3], rays[:, 3], rays[:, 4:7], rays[:, 7] = a["o"], a["mint"], a["d"], a["maxt"]
    cam = np.asarray(s["to_world"], np.float64)[0:3, 3]
    k = n // 4                                     # a quarter of the rays fly at the aperture (radius 0.12) from inside the scene
    src = rng.uniform([-2, 0.2, -1.5], [2, 2.0, 1.5], (k, 3))
    dst = cam + rng.normal(size=(k, 3)) * 0.1
    d = dst - src
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays[:k, 0:3], rays[:k, 4:7] = src, d
    rays[k:k + 500, 7] = rng.uniform(0.05, 2.0, 500)     # bounded segments
    ip, op = os.path.join(td, "all_in.bin"), os.path.join(td, "all_out.bin")
    rays.tofile(ip)
    r = subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", "all", xml, ip, op], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0 or not os.path.exists(op):
        raise RuntimeError(r.stdout[-3000:])
    print([l for l in r.stdout.splitlines() if "special shapes" in l])
    out = np.fromfile(op, np.float32).reshape(n, 4)
    shape = out[:, 1].copy().view(np.int32)
    np.savez_compressed(os.path.join(GOLD, "all.npz"), rays=rays, t=out[:, 0], shape=shape, any=out[:, 2].astype(np.uint8))
    print("all:", n, "rays; hits on kd shapes", (shape >= 0).sum(), "aperture disk", (shape == -2).sum(), "environment sphere", (shape == -3).sum(), "misses", (shape == -1).sum())


def fog_cases(td):
    """SimpleVolumetricPathTracer (src/integrators/path/volpath_simple.cpp) over homogeneous media: index-matched and
    dielectric medium boundaries, isotropic and Henyey-Greenstein phase functions; with the camera outside every medium
    (depth 6 and the short depth 3, where maxInteractions of Scene::evalTransmittance bites) and inside a haze (unbounded)."""
    li_case(td, "fog", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=37))
    li_case(td, "fog_short", S.fog_scene(48, 32, spp=4, max_depth=3, sampler="ctr", seed=41))
    li_case(td, "fog_haze", S.fog_scene(48, 32, spp=4, max_depth=-1, sampler="ctr", seed=43, haze=True))
    # VolumetricPathTracer (src/integrators/path/volpath.cpp): the same scenes with multiple importance sampling
    li_case(td, "fogmis", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=37, kind="volpath"))
    li_case(td, "fogmis_short", S.fog_scene(48, 32, spp=4, max_depth=3, sampler="ctr", seed=41, kind="volpath"))
    li_case(td, "fogmis_haze", S.fog_scene(48, 32, spp=4, max_depth=-1, sampler="ctr", seed=43, haze=True, kind="volpath"))
    # an area light on an analytic sphere inside the haze: the kernels compiled WITH the analytic-light / texture / envmap branches
    li_case(td, "fog_sphere", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=47, haze=True, sphere_light=True))
    li_case(td, "fogmis_sphere", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=53, haze=True, sphere_light=True, kind="volpath"))
    # strictNormals + hideEmitters in the volumetric integrators (volpath_simple.cpp:176-197,223,266; volpath.cpp likewise)
    for name, kind, seed in (("fog_strict", "volpath_simple", 67), ("fogmis_strict", "volpath", 71)):
        sd = S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=seed, haze=True, kind=kind)
        sd.set_integrator(kind, max_depth=6, rr_depth=3, strict_normals=True, hide_emitters=True)
        li_case(td, name, sd)
    # a plastic ball (the kernels outside the LEAN set, independent sampler) seen through a thinlens sensor inside the haze
    PLASTIC = dict(type="plastic", diffuse_reflectance=(0.6, 0.3, 0.2), nonlinear=True)
    li_case(td, "fog_plastic_thinlens", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=73, haze=True, ball_bsdf=PLASTIC, aperture_radius=0.1))
    li_case(td, "fogmis_plastic_thinlens", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=79, haze=True, ball_bsdf=PLASTIC, aperture_radius=0.1, kind="volpath"))
    # the environment map behind the media (pyramid: tests/golden/envmap_pyramid.npz, as envmap_cases dumps it)
    g = np.load(os.path.join(GOLD, "envmap_pyramid.npz"))
    pyramid = [g["level%d" % l].astype(np.float32) for l in range(len(g.files))]
    li_case(td, "fog_envmap", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=59, haze=True, envmap=True, envmap_pyramid=pyramid))
    li_case(td, "fogmis_envmap", S.fog_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=61, haze=True, envmap=True, envmap_pyramid=pyramid, kind="volpath"))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "ldctr":            # round 2
        with tempfile.TemporaryDirectory() as td:
            ldctr_cases(td)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "sobol":            # round 2
        with tempfile.TemporaryDirectory() as td:
            sobol_cases(td)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "texture":          # round 2
        with tempfile.TemporaryDirectory() as td:
            texture_cases(td)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "envmap":           # round 2
        with tempfile.TemporaryDirectory() as td:
            envmap_cases(td)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "analytic_lights":  # round 2
        with tempfile.TemporaryDirectory() as td:
            analytic_lights_cases(td)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "all":              # rayIntersectAll (round 2)
        with tempfile.TemporaryDirectory() as td:
            all_case(td)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "roughplastic":     # round 2
        with tempfile.TemporaryDirectory() as td:
            roughplastic_cases(td)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "fog":              # participating media (round 2)
        with tempfile.TemporaryDirectory() as td:
            fog_cases(td)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "analytic":         # analytic sphere / disk shapes (round 2)
        with tempfile.TemporaryDirectory() as td:
            li_case(td, "analytic", S.analytic_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=31))
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "direct_fan":       # only the cases added in round 2
        with tempfile.TemporaryDirectory() as td:
            for n_em, n_bsdf in ((4, 2), (1, 3), (3, 1), (2, 0)):
                direct_fan_case(td, n_em, n_bsdf)
        sys.exit(0)
    with tempfile.TemporaryDirectory() as td:
        bsdf_case(td)
        emitter_case(td)
        li_case(td, "cornell", S.cornell_box(32, 32, spp=4, max_depth=5, sampler="ctr", seed=7))
        li_case(td, "materials", S.materials_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=11))
        li_case(td, "materials_deep", S.materials_scene(24, 16, spp=2, max_depth=-1, sampler="ctr", seed=3))
        direct_case(td)
        thinlens_case(td)
        strict_case(td)
        for n_em, n_bsdf in ((4, 2), (1, 3), (3, 1), (2, 0)):
            direct_fan_case(td, n_em, n_bsdf)
        li_case(td, "analytic", S.analytic_scene(48, 32, spp=4, max_depth=6, sampler="ctr", seed=31))
        fog_cases(td)
        roughplastic_cases(td)
        all_case(td)
        analytic_lights_cases(td)
        envmap_cases(td)
        texture_cases(td)
        sobol_cases(td)
