"""Golden vectors for BASELINE configs 3/4/5 (the 10 485 762-triangle sphere field), produced by the UNMODIFIED
reference (oracle/_ref) on the CPU of the build container.  The SAH kd-tree build over 10.5 M triangles takes minutes,
which is why this runs once, here, and its outputs are committed:

    python tests/golden/make_golden_c3.py [--rays 131072] [--spp 4096] [--threads 8]

Outputs
  tests/golden/hits_c3.npz   hit records of ShapeKDTree::rayIntersect(ray, its) / rayIntersect(ray) for four ray sets
                             (chord, primary, secondary, shadow).  The rays themselves are NOT stored: they are
                             regenerated bit for bit by c3_rays() below from seeds and from the stored reference `t` of
                             the primary set, so the file stays small.
  tests/golden/img_c3.npz    the reference's own render of config 3 at 240x135 and --spp samples per pixel (classic
                             scheduler, stock `path` plugin, independent sampler, Gaussian filter, maxDepth 8) — the
                             ground truth of the relMSE-vs-spp test.
"""
import argparse
import hashlib
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as g  # noqa: E402

IMG_W, IMG_H = 240, 135


def c3_camera():
    return (0.0, 5.5, -8 * 1.1 - 2), (0.0, 0.3, 0.0)     # scenes.sphere_field's look_at for ny = 8


def c3_primary(S, n):
    """Set 0 (chord rays through the field's bounding sphere) and set 1 (pinhole rays of the config-3 camera)."""
    res = int(round(np.sqrt(n)))
    o, tgt = c3_camera()
    return S.chord_rays(n, center=(0, 0.5, 0), radius=9.0, seed=42, mint=0.0), S.pinhole_rays(res, origin=o, target=tgt, fov_deg=45.0)


def c3_derived(S, r1, t1):
    """Sets 2 and 3, from the primary set and the reference's `t` for it: secondary rays leave the primary hit points
    (mint = Epsilon, as every ray MIPathTracer spawns) in seeded uniform directions folded to the side the camera ray
    came from; shadow segments go from the same points to the scene's light quad (scene.cpp:871-873 conventions)."""
    hit = np.isfinite(t1)
    p = (r1["o"][hit].astype(np.float64) + r1["d"][hit].astype(np.float64) * t1[hit].astype(np.float64)[:, None]).astype(np.float32)
    rng = np.random.Generator(np.random.Philox(77))
    w = S._uniform_sphere(rng.random((p.shape[0], 2), dtype=np.float32).astype(np.float64))
    flip = (w * r1["d"][hit].astype(np.float64)).sum(axis=1) > 0
    w[flip] = -w[flip]
    r2 = np.zeros(p.shape[0], r1.dtype)
    r2["o"], r2["d"] = p, w.astype(np.float32)
    r2["mint"], r2["maxt"] = np.float32(S.EPSILON), np.float32(np.inf)
    r3 = S.shadow_rays(p, light_center=(0, 6.0, 0), light_half=2.0, seed=43)
    return r2, r3


def scene_digest(sd):
    h = hashlib.sha1()
    for s in sd.shapes:
        if s["type"] == "mesh":
            h.update(np.ascontiguousarray(s["V"], np.float32).tobytes())
            h.update(np.ascontiguousarray(s["F"], np.uint32).tobytes())
    return h.hexdigest()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=1 << 17)
    ap.add_argument("--spp", type=int, default=4096)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--skip-image", action="store_true")
    ap.add_argument("--skip-hits", action="store_true")
    ap.add_argument("--curve-only", action="store_true", help="keep the existing reference image, (re)compute only the CPU relMSE-vs-spp curve")
    args = ap.parse_args()
    import oracle as O
    pkg = g.load_package()
    S = pkg.scenes
    out_dir = os.path.join(ROOT, "tests", "golden")
    t0 = time.time()
    sd = S.sphere_field()
    digest = scene_digest(sd)
    print("scene built in %.1f s, sha1 %s" % (time.time() - t0, digest), flush=True)
    with tempfile.TemporaryDirectory() as td:
        xml = sd.to_xml(td, "c3")
        if not args.skip_hits:
            r0, r1 = c3_primary(S, args.rays)
            first, tm1 = O.ref_trace(xml, np.concatenate([r0, r1]), threads=args.threads, timeout=36000)
            print("primary sets traced:", tm1, flush=True)
            t1 = first[r0.shape[0]:, 0]
            r2, r3 = c3_derived(S, r1, t1)
            second, tm2 = O.ref_trace(xml, np.concatenate([r2, r3]), threads=args.threads, timeout=36000)
            print("derived sets traced:", tm2, flush=True)
            out = np.concatenate([first, second])
            sizes = np.array([r0.shape[0], r1.shape[0], r2.shape[0], r3.shape[0]], np.int64)
            shape = out[:, 3].copy().view(np.int32)
            prim = out[:, 4].copy().view(np.int32)
            # global primitive index in ShapeKDTree::build order (m_shapeMap prefix sums; the rectangle counts 1)
            offs = np.cumsum([0] + [s["F"].shape[0] if s["type"] == "mesh" else 1 for s in sd.shapes])
            is_rect = np.array([s["type"] != "mesh" for s in sd.shapes])
            gprim = np.where(shape >= 0, offs[np.maximum(shape, 0)] + np.where(is_rect[np.maximum(shape, 0)], 0, np.maximum(prim, 0)), -1).astype(np.int32)
            np.savez_compressed(os.path.join(out_dir, "hits_c3.npz"), sizes=sizes, n=np.int64(args.rays), digest=np.array(digest),
                                ref_t=out[:, 0].copy(), ref_u=out[:, 1].copy(), ref_v=out[:, 2].copy(), ref_prim=gprim,
                                ref_any=np.packbits(out[:, 14].copy().view(np.int32).astype(np.uint8)),
                                kd_build_ms=np.float64(tm1.get("build_ms") or 0),
                                cpu_closest_mrays=np.float64(tm1.get("closest_mrays") or 0), cpu_threads=np.int64(args.threads))
            print("hits_c3.npz: %d rays, hit rate per set %s" % (out.shape[0], [float(np.isfinite(out[a:b, 0]).mean()) for a, b in zip(np.cumsum(sizes) - sizes, np.cumsum(sizes))]), flush=True)
        if args.curve_only:
            args.skip_image = True
        if not args.skip_image:
            sdi = S.sphere_field(width=IMG_W, height=IMG_H, spp=args.spp)
            xmli = sdi.to_xml(os.path.join(td, "img"), "c3img")
            pfm = os.path.join(td, "c3img.pfm")
            wall, render_s, log = O.ref_render(xmli, pfm, threads=args.threads, classic=True, timeout=36000)
            img = O.read_pfm(pfm)
            np.savez_compressed(os.path.join(out_dir, "img_c3.npz"), img=img.astype(np.float32), spp=np.int64(args.spp), digest=np.array(digest),
                                render_s=np.float64(render_s or 0), wall_s=np.float64(wall), cpu_threads=np.int64(args.threads))
            print("img_c3.npz: %dx%d at %d spp, render %.1f s (wall %.1f s), mean %.4f" % (IMG_W, IMG_H, args.spp, render_s or -1, wall, img.mean()), flush=True)
        # the reference's OWN error curve against that image: relMSE of its renders at 4 / 16 / 64 / 256 spp — what the GPU
        # estimator's curve has to overlap if it is unbiased and has the same variance
        path = os.path.join(out_dir, "img_c3.npz")
        if os.path.exists(path) and not args.skip_hits or args.curve_only:
            z = dict(np.load(path))
            ref = z["img"].astype(np.float64)
            spps, rel, secs = [4, 16, 64, 256], [], []
            for spp in spps:
                sdi = S.sphere_field(width=IMG_W, height=IMG_H, spp=spp)
                xmli = sdi.to_xml(os.path.join(td, "img%d" % spp), "c3img")
                pfm = os.path.join(td, "c3img%d.pfm" % spp)
                wall, render_s, log = O.ref_render(xmli, pfm, threads=args.threads, classic=True, timeout=36000)
                img = O.read_pfm(pfm).astype(np.float64)
                rel.append(float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2))))
                secs.append(float(render_s or 0))
                print("cpu curve: %d spp relMSE %.5g (render %.2f s)" % (spp, rel[-1], secs[-1]), flush=True)
            z.update(cpu_spp=np.array(spps, np.int64), cpu_relmse=np.array(rel), cpu_render_s=np.array(secs))
            np.savez_compressed(path, **z)


if __name__ == "__main__":
    main()
