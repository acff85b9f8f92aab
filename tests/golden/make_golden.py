"""Generates the committed golden vectors by running the UNMODIFIED reference (oracle/_ref, built by
oracle/Makefile.ref from /root/reference) through oracle/raydump.cpp.  Run here, in the build container:

    python tests/golden/make_golden.py

Outputs tests/golden/hits_*.npz: scene arrays + ray batches + the reference's hit records
(ShapeKDTree::rayIntersect(ray, its) and ShapeKDTree::rayIntersect(ray)).
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
S, B = pkg.scenes, pkg.binding
REF = os.path.join(ROOT, "oracle", "_ref")


def ref_trace(xml, rays, threads=4):
    """Runs mtsutil raydump trace and returns the N x 16 float record array."""
    with tempfile.TemporaryDirectory() as td:
        rp, op = os.path.join(td, "rays.bin"), os.path.join(td, "out.bin")
        rays.view(np.float32).tofile(rp)
        r = subprocess.run([os.path.join(REF, "mtsutil"), "raydump", "trace", xml, rp, op, str(threads)],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0 or not os.path.exists(op):
            raise RuntimeError(r.stdout[-3000:])
        out = np.fromfile(op, np.float32).reshape(-1, 16)
    return out


def pack(out):
    return dict(ref_t=out[:, 0].copy(), ref_u=out[:, 1].copy(), ref_v=out[:, 2].copy(),
                ref_shape=out[:, 3].copy().view(np.int32), ref_prim=out[:, 4].copy().view(np.int32),
                ref_p=out[:, 5:8].copy(), ref_ng=out[:, 8:11].copy(), ref_ns=out[:, 11:14].copy(),
                ref_any=out[:, 14].copy().view(np.int32))


def sphere_case(td, n=1536):
    V, N, F = S.displaced_sphere(3, seed=1)
    sd = S.SceneDesc()
    sd.add_mesh(V, N, F)
    sd.add_point_light((0, 5, 0), (1, 1, 1))   # a scene without emitters makes Scene::initialize load `sunsky`
    xml = sd.to_xml(td, "sphere")
    r0 = S.chord_rays(n, radius=1.1, seed=42, mint=0.0)
    r1 = S.chord_rays(n, radius=1.1, seed=7, mint=S.EPSILON)
    r2 = S.pinhole_rays(32, origin=(0, 0.3, -3.0))
    first = ref_trace(xml, r2)
    hitp = first[np.isfinite(first[:, 0])][:, 5:8]
    r3 = S.shadow_rays(hitp, light_center=(0, 3.0, 0.5), light_half=0.2)
    rays = np.concatenate([r0, r1, r2, r3])
    out = ref_trace(xml, rays)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "hits_sphere.npz"), V=V, N=N, F=F,
                        rays=rays.view(np.float32).reshape(-1, 8), **pack(out))
    print("sphere:", rays.shape[0], "rays, hit rate", np.isfinite(out[:, 0]).mean(), "any rate", pack(out)["ref_any"].mean())


def cornell_case(td, n=1536):
    sd = S.cornell_box(64, 64, 1)
    xml = sd.to_xml(td, "cbox")
    r0 = S.chord_rays(n, center=(0, 1, 0), radius=0.95, seed=3, mint=S.EPSILON)
    cam = S.look_at((0, 1, 3.9), (0, 1, 0), (0, 1, 0))
    r1 = S.pinhole_rays(40, origin=(0, 1, 3.9), target=(0, 1, 0), fov_deg=39.3)
    first = ref_trace(xml, r1)
    hitp = first[np.isfinite(first[:, 0])][:, 5:8]
    r2 = S.shadow_rays(hitp, light_center=(0, 1.98, 0), light_half=0.25)
    rays = np.concatenate([r0, r1, r2])
    out = ref_trace(xml, rays)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "hits_cornell.npz"),
                        rays=rays.view(np.float32).reshape(-1, 8), **pack(out))
    print("cornell:", rays.shape[0], "rays, hit rate", np.isfinite(out[:, 0]).mean(), "any rate", pack(out)["ref_any"].mean())


def analytic_case(td, n=2048):
    """Analytic `sphere` / `disk` shapes among meshes (scenes.analytic_scene): chord rays, primary rays, shadow segments."""
    sd = S.analytic_scene(48, 32, 1)
    xml = sd.to_xml(td, "analytic")
    r0 = S.chord_rays(n, center=(0.2, 0.7, 0.3), radius=2.4, seed=5, mint=S.EPSILON)
    r1 = S.pinhole_rays(48, origin=(0.3, 3.0, -4.5), target=(0, 0.4, 0), fov_deg=40.0)
    first = ref_trace(xml, r1)
    hitp = first[np.isfinite(first[:, 0])][:, 5:8]
    r2 = S.shadow_rays(hitp, light_center=(0.5, 2.6, 0.3), light_half=0.3)
    rays = np.concatenate([r0, r1, r2])
    out = ref_trace(xml, rays)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "hits_analytic.npz"),
                        rays=rays.view(np.float32).reshape(-1, 8), **pack(out))
    sh = pack(out)["ref_shape"]
    print("analytic:", rays.shape[0], "rays, hit rate", np.isfinite(out[:, 0]).mean(), "hits per shape", np.bincount(sh[sh >= 0]))


def light_case(td):
    """ShapeKDTree::rayIntersect(ray, t, shape, n, uv) (skdtree.cpp:144-204) on the ray sets of the two files above plus a
    mesh WITH texture coordinates -> light.npz (raydump `light`: t, shape, n, uv)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers
    res = {}
    for kind in ("sphere", "cornell", "uvsphere"):
        gold = np.load(os.path.join(ROOT, "tests", "golden", "hits_%s.npz" % ("sphere" if kind == "uvsphere" else kind)))
        if kind == "uvsphere":
            sd = helpers.uv_sphere_desc(S, gold)
        else:
            sd = helpers.desc_from_golden(S, gold, kind)
        xml = sd.to_xml(td, "light_" + kind)
        rays = gold["rays"]
        rp, op = os.path.join(td, "rays.bin"), os.path.join(td, "out.bin")
        np.ascontiguousarray(rays, np.float32).tofile(rp)
        r = subprocess.run([os.path.join(REF, "mtsutil"), "raydump", "light", xml, rp, op], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0 or not os.path.exists(op):
            raise RuntimeError(r.stdout[-3000:])
        out = np.fromfile(op, np.float32).reshape(-1, 8)
        res[kind] = out
        print("light/%s: hit rate %.3f, mean |uv| %.3f" % (kind, np.isfinite(out[:, 0]).mean(), np.abs(out[:, 5:7]).mean()))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "light.npz"), **res)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "analytic":
        with tempfile.TemporaryDirectory() as td:
            analytic_case(td)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "light":
        with tempfile.TemporaryDirectory() as td:
            light_case(td)
        sys.exit(0)
    with tempfile.TemporaryDirectory() as td:
        sphere_case(td)
        cornell_case(td)
        light_case(td)
        analytic_case(td)
