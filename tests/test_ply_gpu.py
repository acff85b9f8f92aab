"""GPU parity tests of the device mesh-file decoders (mtsb_scene_add_ply / _add_serialized, csrc/ply.cu) against what the unmodified
reference's `ply` / `serialized` plugins + TriMesh::configure produce for the same file images: committed golden meshes (tests/golden/ply_meshes.npz,
made by tests/golden/make_golden_ply.py) and, where oracle/_ref is present, the live reference.

Bar: triangle indices, texture coordinates and untransformed positions bit-exact; transformed positions and stored normals
within 2e-6 (the reference is a -ffast-math build: its matrix product may reassociate); computed vertex normals within
2e-5 (float atomics add the face terms in another order than the reference's serial loop, and asinf differs by an ulp)."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

import oracle as O
import ply_cases

pytestmark = pytest.mark.gpu


def _decode(pkg, ctx, case, extra_first=None, kind="ply"):
    B = pkg.binding
    sc = B.Scene(ctx)
    bs = sc.add_bsdf(B.BSDF_DIFFUSE, (0.5, 0.5, 0.5))
    if extra_first is not None:
        sc.add_mesh(extra_first[0], extra_first[1], bsdf=bs)
    add = sc.add_ply if kind == "ply" else sc.add_serialized
    shape = add(case["image"], bsdf=bs, **case["kwargs"])
    sc.commit(B.BUILDER_LBVH)
    return sc, shape


def _check_against(got, ref, name, computed, xform):
    P, N, UV, T = got
    rP, rN, rUV, rT = ref
    assert P.shape == rP.shape and T.shape == rT.shape, name
    assert (T == rT).all(), name
    if xform:
        assert np.abs(P - rP).max() <= 2e-6 * max(1.0, np.abs(rP).max()), name
    else:
        assert (P.view(np.uint32) == rP.view(np.uint32)).all(), name
    if rUV is not None:
        assert (UV.view(np.uint32) == rUV.view(np.uint32)).all(), name
    if rN is not None:
        assert np.abs(N - rN).max() <= (2e-5 if computed else 2e-6), (name, np.abs(N - rN).max())


def _golden_ref(gold, name):
    g = lambda k: gold[name + k] if (name + k) in gold.files else None
    return g("_P"), g("_N"), g("_UV"), g("_T")


def _all_cases(pkg):
    out = {n: (c, "ply") for n, c in ply_cases.cases().items()}
    out.update({n: (c, "serialized") for n, c in ply_cases.ser_cases(pkg.scenes).items()})
    return out


CASE_NAMES = sorted(ply_cases.cases().keys()) + ["ser_v3_mirror_xform", "ser_v4_double_computed_xform", "ser_v4_double_flip", "ser_v4_facenormals_flip", "ser_v4_n_uv"]


@pytest.mark.parametrize("name", CASE_NAMES)
def test_device_ply_decode_matches_reference_loader(pkg, ctx, golden, name):
    case, kind = _all_cases(pkg)[name]
    gold = golden("ply_meshes.npz")
    sc, shape = _decode(pkg, ctx, case, kind=kind)
    got = sc.read_mesh(shape)
    ref = _golden_ref(gold, name)
    _check_against(got, ref, name, computed="computed" in name, xform="xform" in name)
    if name in ("ser_v4_n_uv", "ser_v4_double_flip"):
        # the `serialized` plugin leaves stored normals alone without a toWorld (serialized.cpp:184-195): bit-exact, flipped or not
        assert (got[1].view(np.uint32) == ref[1].view(np.uint32)).all()
    if ref[1] is None:      # faceNormals: the shape record says "no vertex normals"
        surf = sc.ray_intersect_surface(np.array([[0, 0, 5, 0, 0, 0, -1, np.inf]], np.float32))
        assert np.isfinite(surf["t"][0]) and np.abs(surf["ns"][0] - surf["ng"][0]).max() < 1e-6
    if os.path.exists(os.path.join(O.REF_DIR, "mtsutil")):        # live reference (travels to the GPU box)
        with tempfile.TemporaryDirectory() as td:
            with open(os.path.join(td, "m." + kind), "wb") as f:
                f.write(case["image"])
            with open(os.path.join(td, "m.xml"), "w") as f:
                f.write(ply_cases.scene_xml("m." + kind, case["xml"], kind))
            r = subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", "mesh", os.path.join(td, "m.xml"), os.path.join(td, "m.bin")],
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            assert r.returncode == 0, r.stdout[-2000:]
            live = ply_cases.read_mesh_dump(os.path.join(td, "m.bin"))[0]
        _check_against(got, live, name + " (live)", computed="computed" in name, xform="xform" in name)


def test_ply_mesh_traces_like_the_host_staged_mesh(pkg, ctx, golden):
    """Same arrays, same structure: a scene holding a host-staged mesh followed by the decoded PLY answers every ray like the
    scene built from the reference's arrays through mtsb_scene_add_mesh (shape offsets and vertex bases included)."""
    B = pkg.binding
    case = ply_cases.cases()["tri_n_uv"]
    gold = golden("ply_meshes.npz")
    P, N, UV, T = _golden_ref(gold, "tri_n_uv")
    floorV = np.array([[-3, -3, -1], [3, -3, -1], [3, 3, -1], [-3, 3, -1]], np.float32)
    floorT = np.array([[0, 1, 2], [0, 2, 3]], np.uint32)
    sc, shape = _decode(pkg, ctx, case, extra_first=(floorV, floorT))
    assert shape == 1
    ref = B.Scene(ctx)
    bs = ref.add_bsdf(B.BSDF_DIFFUSE, (0.5, 0.5, 0.5))
    ref.add_mesh(floorV, floorT, bsdf=bs)
    ref.add_mesh(P, T, normals=N, texcoords=UV, bsdf=bs)
    ref.commit(B.BUILDER_LBVH)
    fP, fN, fUV, fT = sc.read_mesh(0, sizes=(4, 2))
    assert (fP == floorV).all() and (fT == floorT).all()
    rng = np.random.default_rng(3)
    n = 20000
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = rng.uniform(-2, 2, (n, 3)) + [0, 0, 3]
    d = rng.normal(size=(n, 3)) - [0, 0, 1.5]
    rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 7] = np.inf
    a, b = sc.ray_intersect_surface(rays), ref.ray_intersect_surface(rays)
    assert np.isfinite(b["t"]).mean() > 0.3
    for k in ("t", "prim", "p", "ng", "uv"):
        assert (a[k].view(np.uint32) == b[k].view(np.uint32)).all(), k
    # stored normals are re-normalised by the loader (ply.cpp:210): the golden normals went through the reference's
    # -ffast-math rsqrt, the decoded ones through an IEEE division
    hit = np.isfinite(b["t"])
    assert np.abs(a["ns"][hit] - b["ns"][hit]).max() <= 2e-6


def test_ply_decoder_refuses_what_it_cannot_decode(pkg, ctx):
    B = pkg.binding
    case = ply_cases.cases()["tri_n_uv"]
    img = case["image"]

    def attempt(image, commit=True):
        sc = B.Scene(ctx)
        bs = sc.add_bsdf(B.BSDF_DIFFUSE, (0.5, 0.5, 0.5))
        sc.add_ply(image, bsdf=bs)
        if commit:
            sc.commit(B.BUILDER_LBVH)

    with pytest.raises(B.MtsbError, match="binary_little_endian"):
        attempt(img.replace(b"binary_little_endian", b"ascii"))
    with pytest.raises(B.MtsbError, match="neither all triangles nor all quads"):
        attempt(img[:-5])                                     # truncated
    with pytest.raises(B.MtsbError, match="no 'ply' magic"):
        attempt(b"plx\n" + img[4:])
    # a 2-gon and a quad take the bytes of two triangles: the sizes pass, the count check on the device does not
    hdr = b"ply\nformat binary_little_endian 1.0\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\nelement face 2\nproperty list uchar int vertex_indices\nend_header\n"
    verts = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], "<f4").tobytes()
    mixed = hdr + verts + bytes([2]) + np.array([0, 1], "<i4").tobytes() + bytes([4]) + np.array([0, 1, 2, 3], "<i4").tobytes()
    with pytest.raises(B.MtsbError, match="mixes triangles and quads") as e:
        attempt(mixed)
    assert e.value.code == B.EUNSUPPORTED
    bad = hdr + verts + bytes([3]) + np.array([0, 1, 2], "<i4").tobytes() + bytes([3]) + np.array([0, 2, 7], "<i4").tobytes()
    with pytest.raises(B.MtsbError, match="references a vertex"):
        attempt(bad)
    # `serialized`: wrong magic, shape index out of range, a stream cut short
    ser = ply_cases.ser_cases(pkg.scenes)["ser_v4_n_uv"]["image"]

    def attempt_ser(image, index=0):
        sc = B.Scene(ctx)
        bs = sc.add_bsdf(B.BSDF_DIFFUSE, (0.5, 0.5, 0.5))
        sc.add_serialized(image, shape_index=index, bsdf=bs)

    assert B.serialized_probe(ser, 1) == (100, 162) and B.ply_probe(img) == (169, 288)
    with pytest.raises(B.MtsbError, match="invalid file format"):
        attempt_ser(b"\x00\x00" + ser[2:])
    with pytest.raises(B.MtsbError, match="out of range"):
        attempt_ser(ser, 3)
    with pytest.raises(B.MtsbError, match="stream"):
        attempt_ser(ser[:2000])
    neg = hdr + verts + bytes([3]) + np.array([0, 1, 2], "<i4").tobytes() + bytes([3]) + np.array([0, 2, -1], "<i4").tobytes()
    with pytest.raises(B.MtsbError, match="references a vertex"):
        attempt(neg)


def test_ply_decode_at_config2_size(pkg, ctx):
    """BASELINE config 2's mesh (1 310 720 triangles) as a binary PLY image: the decoded positions and indices equal the arrays the
    image was written from bit for bit, stored normals to the last bit of their re-normalisation, and the scene built from the
    image answers 2 M chord rays exactly like the scene built from the arrays (same arrays -> same BVH -> same hits)."""
    import bench
    S, B = pkg.scenes, pkg.binding
    sd = bench.build_mesh(S)
    m = sd.shapes[0]
    image = S.ply_image(m["V"], m["N"], m["F"])
    assert B.ply_probe(image) == (m["V"].shape[0], m["F"].shape[0])
    sc_ply = sd.upload(ctx, B.BUILDER_LBVH, ply_images={0: image})
    sc_arr = sd.upload(ctx, B.BUILDER_LBVH)
    P, N, UV, T = sc_ply.read_mesh(0)
    assert (P.view(np.uint32) == np.asarray(m["V"], np.float32).view(np.uint32)).all()
    assert (T == np.asarray(m["F"], np.uint32)).all()
    assert np.abs(N - m["N"]).max() <= 2e-7
    assert sc_ply.build_stats()["ply_decode_ms"] > 0 and sc_arr.build_stats()["ply_decode_ms"] == 0
    rays = S.chord_rays(1 << 21, radius=1.1, seed=5)
    a, b = sc_ply.ray_intersect(rays), sc_arr.ray_intersect(rays)
    assert np.isfinite(b["t"]).mean() > 0.5
    for k in ("t", "u", "v", "prim"):
        assert (a[k].view(np.uint32) == b[k].view(np.uint32)).all(), k
    sc_ply.close(); sc_arr.close()
