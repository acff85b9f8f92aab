"""Host logic of the scene description (mitsuba-im_b200/scenes.py) on the CPU: what goes into the C ABI's BSDF / camera
records and what goes into the reference's XML must describe the same plugin — the GPU parity tests rely on it."""
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_enum(prefix):
    text = open(os.path.join(ROOT, "include", "mtsb200.h")).read()
    return {m.group(1): int(m.group(2), 0) for m in re.finditer(r"(%s\w+)\s*=\s*(0x[0-9a-fA-F]+|\d+)" % prefix, text)}


def test_binding_constants_match_the_header(pkg):
    B = pkg.binding
    enum = _header_enum("MTSB_BSDF_")
    for name, value in enum.items():
        assert getattr(B, name[len("MTSB_"):]) == value, name
    assert len(enum) >= 8


def test_every_material_packs_into_the_abi_record(pkg):
    S, B = pkg.scenes, pkg.binding
    enum = _header_enum("MTSB_BSDF_")
    plugin_of = {enum["MTSB_BSDF_DIFFUSE"]: "diffuse", enum["MTSB_BSDF_DIELECTRIC"]: "dielectric", enum["MTSB_BSDF_PLASTIC"]: "plastic",
                 enum["MTSB_BSDF_ROUGHCONDUCTOR"]: "roughconductor", enum["MTSB_BSDF_CONDUCTOR"]: "conductor",
                 enum["MTSB_BSDF_ROUGHDIELECTRIC"]: "roughdielectric"}
    seen = set()
    for m in S.MATERIALS:
        tid, params = S.SceneDesc._bsdf_params(m)
        base, two = tid & 0xff, bool(tid & B.BSDF_TWOSIDED_FLAG)
        assert base in plugin_of and len(params) <= 15 and np.isfinite(params).all(), m
        assert two == bool(m.get("twosided"))
        xml = S.SceneDesc._bsdf_xml(m)
        assert 'type="%s"' % plugin_of[base] in xml, (m, xml)
        assert ('type="twosided"' in xml) == two
        seen.add(base)
    assert seen == set(plugin_of), "MATERIALS must exercise every BSDF of the GPU subset"


def test_relative_ior_and_conductor_scaling(pkg):
    S = pkg.scenes
    _, p = S.SceneDesc._bsdf_params(dict(type="dielectric", int_ior=1.5, ext_ior=1.25))
    assert abs(p[0] - 1.2) < 1e-12                                   # eta = intIOR / extIOR (dielectric.cpp:150-165)
    _, p = S.SceneDesc._bsdf_params(dict(type="conductor", eta=(2, 4, 6), k=(1, 2, 3), ext_eta=2.0))
    assert p[:6] == [1, 2, 3, 0.5, 1, 1.5]                           # eta / extEta, k / extEta (conductor.cpp:174-177)
    _, p = S.SceneDesc._bsdf_params(dict(type="roughdielectric", alpha_u=0.1, alpha_v=0.3, distribution="ggx", sample_visible=False))
    assert p[0:2] == [0.1, 0.3] and p[11] == 1 and p[12] == 0.0


def test_thinlens_sensor_description(pkg, tmp_path):
    S = pkg.scenes
    sd = S.cornell_box(16, 16, spp=1, max_depth=2)
    cam = sd.camera()
    assert cam.aperture_radius == 0.0                                # perspective: no aperture sample is drawn
    s = sd.sensor
    sd.set_sensor(s["to_world"], s["fov"], s["width"], s["height"], aperture_radius=0.05, focus_distance=3.5)
    cam = sd.camera()
    assert abs(cam.aperture_radius - 0.05) < 1e-7 and abs(cam.focus_distance - 3.5) < 1e-7
    xml = open(sd.to_xml(str(tmp_path), "tl")).read()
    assert '<sensor type="thinlens">' in xml and 'name="apertureRadius" value="0.05"' in xml and 'name="focusDistance" value="3.5"' in xml


def test_rough_transmittance_slices_of_roughplastic(pkg):
    """hostmath.rough_transmittance_slices restates RoughTransmittance::setEta / setAlpha / evalDiffuse (src/bsdfs/rtrans.h:
    243-370) on the reference's own data/microfacet tables: smooth limit and packing into the 15-float ABI record."""
    import importlib
    import oracle as O
    H = importlib.import_module(pkg.__name__ + ".hostmath")
    S, B = pkg.scenes, pkg.binding
    d = os.path.join(O.REF_DIR, "data", "microfacet")
    if not os.path.exists(os.path.join(d, "beckmann.dat")):
        import pytest
        pytest.fail("oracle/_ref/data/microfacet is missing: run __graft_entry__.build()")
    eta = 1.49 / 1.000277
    for dist in ("beckmann", "ggx"):
        t = H.load_rough_transmittance(os.path.join(d, dist + ".dat"))
        assert t["trans"].shape == (100, 50, 100) and t["diff"].shape == (100, 50)
        ext, fdr = H.rough_transmittance_slices(t, eta, 0.01)
        assert ext.shape == (100,) and ext.dtype == np.float32
        # near-smooth interface: the transmittance at normal incidence is 1 - F0, the internal diffuse reflectance that of `plastic`
        f0 = ((eta - 1) / (eta + 1)) ** 2
        assert abs(ext[-1] - (1 - f0)) < 2e-3
        assert abs(fdr - H.fresnel_diffuse_reflectance(1 / eta)) < 0.01
    S.MICROFACET_DIR = d
    for m in S.ROUGH_PLASTICS:
        m = dict(m)
        tid, params = S.SceneDesc._bsdf_params(m)
        assert (tid & 0xff) == B.BSDF_ROUGHPLASTIC and len(params) == 13 and np.isfinite(params).all()
        assert m["_table"].shape == (100,) and 'type="roughplastic"' in S.SceneDesc._bsdf_xml(m)


def test_ld_restatement_matches_the_reference_qmc_functions(golden):
    """tests/ld_restate.py radical_inverse2 / sobol2 against radicalInverse2Single / sobol2Single of the reference
    (tests/golden/qmc.npz, oracle/raydump.cpp `qmc`), bit for bit; and the hashed permutation is one."""
    import ld_restate as L
    g = golden("qmc.npz")
    assert (L.radical_inverse2(g["n"], g["s0"]).view(np.uint32) == g["radical_inverse"].view(np.uint32)).all()
    assert (L.sobol2(g["n"], g["s1"]).view(np.uint32) == g["sobol2"].view(np.uint32)).all()
    for log2n in (0, 1, 4, 6, 10):
        n = 1 << log2n
        for p in (0, 1, 0x12345678, 0xffffffff):
            assert sorted(L.permute(np.arange(n), n - 1, p).tolist()) == list(range(n))
