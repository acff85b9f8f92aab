"""Procedural scenes and ray batches for the five BASELINE.json configs, described ONCE and emitted twice:
as Mitsuba XML + binary PLY for the reference (oracle/_ref) and as flat arrays for the C ABI (binding.Scene).

Host-side helpers only (numpy); nothing here traces a ray.  The camera / film / filter set-up restates
PerspectiveCameraImpl::configure (src/sensors/perspective.cpp:126-180), Transform::lookAt / perspective
(src/libcore/transform.cpp:99-123,191-214) and ReconstructionFilter::configure (src/libcore/rfilter.cpp:37-55)
so that both back ends see the same sensor.
"""
import math
import os

import numpy as np

from . import binding as B

EPSILON = 1e-4
# the renderer's data/microfacet directory (rough-transmittance tables read by `roughplastic`, src/bsdfs/rtrans.h:95-97)
MICROFACET_DIR = os.environ.get("MTSB_MICROFACET_DIR")
# the direction numbers of the `sobol` sampler as uint32 words (include/mtsb200.h: mtsb_render_set_sobol); set by the host
SOBOL_TABLES = None
SHADOW_EPSILON = 1e-3


# ------------------------------------------------------------------------------------------------ meshes
def icosphere(subdiv):
    """Unit icosphere: 20 * 4**subdiv triangles, shared vertices."""
    t = (1.0 + math.sqrt(5.0)) / 2.0
    V = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    V /= np.linalg.norm(V, axis=1, keepdims=True)
    F = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                  [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11],
                  [6, 2, 10], [8, 6, 7], [9, 8, 1]], dtype=np.int64)
    for _ in range(subdiv):
        e = np.concatenate([F[:, [0, 1]], F[:, [1, 2]], F[:, [2, 0]]], axis=0)
        es = np.sort(e, axis=1)
        key = es[:, 0] * (V.shape[0] + 1) + es[:, 1]
        uniq, inv = np.unique(key, return_inverse=True)
        a = uniq // (V.shape[0] + 1)
        b = uniq % (V.shape[0] + 1)
        mid = V[a] + V[b]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        base = V.shape[0]
        V = np.concatenate([V, mid], axis=0)
        n = F.shape[0]
        m01, m12, m20 = base + inv[:n], base + inv[n:2 * n], base + inv[2 * n:]
        F = np.concatenate([np.stack([F[:, 0], m01, m20], 1), np.stack([F[:, 1], m12, m01], 1),
                            np.stack([F[:, 2], m20, m12], 1), np.stack([m01, m12, m20], 1)], axis=0)
    return V, F


def fbm(P, seed, octaves=4):
    """Deterministic smooth pseudo-noise in [-1,1]: octaves of randomly oriented sinusoids (seeded)."""
    rng = np.random.Generator(np.random.Philox(seed))
    out = np.zeros(P.shape[0])
    amp, freq, norm = 1.0, 2.0, 0.0
    for _ in range(octaves):
        acc = np.zeros(P.shape[0])
        for _ in range(3):
            d = rng.normal(size=3)
            d /= np.linalg.norm(d)
            acc += np.sin(freq * (P @ d) + rng.uniform(0, 2 * math.pi))
        out += amp * acc / 3.0
        norm += amp
        amp *= 0.5
        freq *= 2.0
    return out / norm


def vertex_normals(V, F):
    fn = np.cross(V[F[:, 1]] - V[F[:, 0]], V[F[:, 2]] - V[F[:, 0]])
    N = np.zeros_like(V)
    for k in range(3):
        np.add.at(N, F[:, k], fn)
    N /= np.maximum(np.linalg.norm(N, axis=1, keepdims=True), 1e-30)
    return N


def displaced_sphere(subdiv, seed=1, radius=1.0, center=(0, 0, 0), amplitude=0.05):
    """Config 2/3 geometry: icosphere displaced radially by amplitude * fbm(seed); returns float32 V, N and uint32 F."""
    V, F = icosphere(subdiv)
    V = V * (1.0 + amplitude * fbm(V, seed))[:, None]
    N = vertex_normals(V, F)
    V = V * radius + np.asarray(center, dtype=np.float64)
    return V.astype(np.float32), N.astype(np.float32), F.astype(np.uint32)


def ply_image(V, N, F, UV=None):
    """Binary little-endian PLY file image with per-vertex normals (so TriMesh::computeNormals keeps them as-is) and optional
    texture coordinates (properties u, v)."""
    V = np.asarray(V, np.float32)
    F = np.asarray(F, np.int32)
    hdr = "ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n" % V.shape[0]
    if N is not None:
        hdr += "property float nx\nproperty float ny\nproperty float nz\n"
    if UV is not None:
        hdr += "property float u\nproperty float v\n"
    hdr += "element face %d\nproperty list uchar int vertex_indices\nend_header\n" % F.shape[0]
    vd = V if N is None else np.concatenate([V, np.asarray(N, np.float32)], axis=1)
    if UV is not None:
        vd = np.concatenate([vd, np.asarray(UV, np.float32)], axis=1)
    fd = np.empty(F.shape[0], dtype=[("n", "u1"), ("i", "<i4", 3)])
    fd["n"] = 3
    fd["i"] = F
    return hdr.encode("ascii") + np.ascontiguousarray(vd, "<f4").tobytes() + fd.tobytes()


def serialized_image(meshes, version=4, double=False, level=6):
    """A `serialized` file image (TriMesh::serialize layout, src/librender/trimesh.cpp): meshes = list of dicts with V, F and
    optionally N, UV, C (colours), name; one deflate stream per mesh behind a 4-byte header, offset dictionary at the end."""
    import struct
    import zlib
    ft = "<f8" if double else "<f4"
    blobs, offsets, pos = [], [], 0
    for m in meshes:
        V, F = np.asarray(m["V"]), np.asarray(m["F"])
        flags = (0x2000 if double else 0x1000) | (1 if m.get("N") is not None else 0) | (2 if m.get("UV") is not None else 0) | (8 if m.get("C") is not None else 0)
        payload = struct.pack("<I", flags)
        if version == 4:
            payload += m.get("name", "mesh").encode("utf-8") + b"\0"
        payload += struct.pack("<QQ", V.shape[0], F.shape[0]) + np.ascontiguousarray(V, ft).tobytes()
        for k in ("N", "UV", "C"):
            if m.get(k) is not None:
                payload += np.ascontiguousarray(m[k], ft).tobytes()
        payload += np.ascontiguousarray(F, "<u4").tobytes()
        blob = struct.pack("<HH", 0x041C, version) + zlib.compress(payload, level)
        offsets.append(pos); pos += len(blob); blobs.append(blob)
    tail = b"".join(struct.pack("<Q" if version == 4 else "<I", o) for o in offsets) + struct.pack("<I", len(offsets))
    return b"".join(blobs) + tail


def write_ply(path, V, N, F, UV=None):
    with open(path, "wb") as f:
        f.write(ply_image(V, N, F, UV))


# ------------------------------------------------------------------------------------------------ transforms
def write_pfm(path, img):
    """Little-endian colour PFM, rows bottom-up (what the reference's Bitmap reads natively)."""
    img = np.asarray(img, np.float32)
    with open(path, "wb") as f:
        f.write(b"PF\n%d %d\n-1.0\n" % (img.shape[1], img.shape[0]))
        f.write(np.ascontiguousarray(img[::-1]).astype("<f4").tobytes())


def sky_image(width=64, height=32, seed=3):
    """A synthetic latitude-longitude radiance map: horizon gradient, a bright warm sun, a dim cool second light, mild noise."""
    v, u = np.meshgrid((np.arange(height) + 0.5) / height, (np.arange(width) + 0.5) / width, indexing="ij")
    theta, phi = v * math.pi, u * 2 * math.pi
    d = np.stack([np.sin(phi) * np.sin(theta), np.cos(theta), -np.cos(phi) * np.sin(theta)], axis=-1)
    img = np.empty((height, width, 3), np.float32)
    up = np.clip(d[..., 1], -1, 1)
    img[...] = (0.25 + 0.35 * np.clip(up, 0, 1))[..., None] * np.array([0.55, 0.7, 1.0]) + (0.12 * np.clip(-up, 0, 1))[..., None] * np.array([0.5, 0.4, 0.3])
    for c, w, col in (((0.45, 0.75, -0.48), 40.0, (30.0, 24.0, 15.0)), ((-0.7, 0.35, 0.62), 12.0, (1.5, 2.5, 4.0))):
        c = np.asarray(c) / np.linalg.norm(c)
        img += (np.exp(w * ((d * c).sum(axis=-1) - 1.0)))[..., None] * np.array(col, np.float32)
    rng = np.random.default_rng(seed)
    img *= (1.0 + 0.05 * rng.standard_normal((height, width, 1))).astype(np.float32)
    return np.maximum(img, 1e-3).astype(np.float32)


def translate(x, y, z):
    M = np.eye(4)
    M[:3, 3] = [x, y, z]
    return M


def scale(x, y=None, z=None):
    y = x if y is None else y
    z = x if z is None else z
    return np.diag([x, y, z, 1.0])


def rotate(axis, angle_deg):
    a = np.asarray(axis, dtype=np.float64)
    a /= np.linalg.norm(a)
    s, c = math.sin(math.radians(angle_deg)), math.cos(math.radians(angle_deg))
    x, y, z = a
    R = np.array([[x * x + (1 - x * x) * c, x * y * (1 - c) - z * s, x * z * (1 - c) + y * s, 0],
                  [x * y * (1 - c) + z * s, y * y + (1 - y * y) * c, y * z * (1 - c) - x * s, 0],
                  [x * z * (1 - c) - y * s, y * z * (1 - c) + x * s, z * z + (1 - z * z) * c, 0],
                  [0, 0, 0, 1]])
    return R


def look_at(origin, target, up):
    p, t, up = (np.asarray(v, dtype=np.float64) for v in (origin, target, up))
    d = (t - p) / np.linalg.norm(t - p)
    left = np.cross(up, d)
    left /= np.linalg.norm(left)
    new_up = np.cross(d, left)
    M = np.eye(4)
    M[:3, 0], M[:3, 1], M[:3, 2], M[:3, 3] = left, new_up, d, p
    return M


def _mat_xml(M):
    return '<matrix value="%s"/>' % " ".join(repr(float(np.float32(v))) for v in np.asarray(M).reshape(-1))


# ------------------------------------------------------------------------------------------------ sensor / film
def filter_table(kind, radius=None, stddev=0.5):
    """(radius, 32-entry table, border) as ReconstructionFilter::configure computes them (float32)."""
    if kind == "box":
        r = np.float32((0.5 if radius is None else radius)) + np.float32(1e-5)
        ev = lambda x: np.float32(1.0) if abs(x) <= r else np.float32(0.0)
    elif kind == "gaussian":
        r = np.float32(4 * stddev)
        alpha = -1.0 / (2.0 * stddev * stddev)
        ev = lambda x: np.float32(max(0.0, math.exp(alpha * x * x) - math.exp(alpha * float(r) * float(r))))
    else:
        raise ValueError(kind)
    vals = np.zeros(32, np.float32)
    s = np.float32(0)
    for i in range(31):
        vals[i] = ev(float(r) * i / 31.0)
        s += vals[i]
    s = s * np.float32(2) * r / np.float32(31)
    vals[:31] *= np.float32(1.0) / s
    return float(r), vals, int(math.ceil(float(r) - 0.5))


def make_camera(to_world, fov_deg, width, height, near=1e-2, far=1e4, fov_axis="x", crop=None):
    aspect = width / float(height)
    if fov_axis == "y":
        fov_deg = math.degrees(2 * math.atan(math.tan(0.5 * math.radians(fov_deg)) * aspect))
    cx, cy, cw, ch = (0, 0, width, height) if crop is None else crop
    rel_size = (cw / width, ch / height)
    rel_off = (cx / width, cy / height)
    recip = 1.0 / (far - near)
    cot = 1.0 / math.tan(math.radians(fov_deg / 2.0))
    persp = np.array([[cot, 0, 0, 0], [0, cot, 0, 0], [0, 0, far * recip, -near * far * recip], [0, 0, 1, 0]], dtype=np.float64)
    cam_to_sample = (scale(1.0 / rel_size[0], 1.0 / rel_size[1], 1.0) @ translate(-rel_off[0], -rel_off[1], 0.0) @
                     scale(-0.5, -0.5 * aspect, 1.0) @ translate(-1.0, -1.0 / aspect, 0.0) @ persp)
    sample_to_cam = np.linalg.inv(cam_to_sample)
    cam = B.Camera()
    cam.sample_to_camera[:] = [float(v) for v in sample_to_cam.reshape(-1)]
    cam.camera_to_world[:] = [float(v) for v in np.asarray(to_world, dtype=np.float64).reshape(-1)]
    cam.near_clip, cam.far_clip = near, far
    cam.inv_resolution[:] = [1.0 / cw, 1.0 / ch]
    return cam


def make_film(width, height, rfilter="box", channels=5, stddev=0.5):
    r, table, border = filter_table(rfilter, stddev=stddev)
    f = B.FilmDesc()
    f.width, f.height = width, height
    f.offset[:] = [0, 0]
    f.border, f.channels, f.filter_radius = border, channels, r
    f.filter_table[:] = [float(v) for v in table]
    f.crop_offset[:] = [0, 0]
    f.crop_size[:] = [width, height]
    return f


def make_integrator(kind="path", max_depth=-1, rr_depth=5, strict_normals=False, hide_emitters=False,
                    emitter_samples=1, bsdf_samples=1):
    d = B.IntegratorDesc()
    d.type = dict(path=B.INTEGRATOR_PATH, direct=B.INTEGRATOR_DIRECT, volpath_simple=B.INTEGRATOR_VOLPATH_SIMPLE, volpath=B.INTEGRATOR_VOLPATH)[kind]
    d.max_depth, d.rr_depth = max_depth, rr_depth
    d.strict_normals, d.hide_emitters = int(strict_normals), int(hide_emitters)
    d.emitter_samples, d.bsdf_samples = emitter_samples, bsdf_samples
    return d


# ------------------------------------------------------------------------------------------------ scene description
class SceneDesc:
    """Back-end neutral scene: shapes (mesh | rectangle) with a BSDF dict and an optional area-emitter radiance,
    plus free emitters (point, constant), a perspective sensor and an integrator."""

    def __init__(self):
        self.shapes = []
        self.emitters = []          # free emitters: dict(type='point'|'constant', ...)
        self.media = []             # homogeneous media: dict(sigma_a, sigma_s, phase='isotropic'|'hg', g, strategy)
        self.sensor = None
        self.integrator = dict(kind="path", max_depth=-1, rr_depth=5)
        self.camera_medium = None   # index into self.media: the medium the sensor sits in (Sensor::getMedium)
        self.sampler = dict(kind="independent", spp=4, seed=0)

    # -- building --------------------------------------------------------------------------------
    def add_mesh(self, V, N, F, bsdf=None, emitter=None, name=None, UV=None):
        self.shapes.append(dict(type="mesh", V=np.asarray(V, np.float32), N=None if N is None else np.asarray(N, np.float32),
                                F=np.asarray(F, np.uint32), UV=None if UV is None else np.asarray(UV, np.float32), bsdf=bsdf or dict(type="diffuse"), emitter=emitter, name=name))
        return len(self.shapes) - 1

    def add_rectangle(self, to_world, bsdf=None, emitter=None, flip_normals=False):
        self.shapes.append(dict(type="rectangle", to_world=np.asarray(to_world, np.float64), bsdf=bsdf or dict(type="diffuse"),
                                emitter=emitter, flip_normals=flip_normals))
        return len(self.shapes) - 1

    def add_sphere(self, center, radius, bsdf=None, flip_normals=False, emitter=None):
        """Analytic `sphere` (src/shapes/sphere.cpp) given by center + radius; emitter = radiance of an area light on it."""
        self.shapes.append(dict(type="sphere", center=tuple(float(x) for x in center), radius=float(radius),
                                bsdf=bsdf if (bsdf is not None or emitter is not None) else dict(type="diffuse"),
                                emitter=emitter, flip_normals=flip_normals))
        return len(self.shapes) - 1

    def add_disk(self, to_world, bsdf=None, flip_normals=False, emitter=None):
        """Analytic `disk` (src/shapes/disk.cpp): the unit disk of the object-space xy plane under `to_world`."""
        self.shapes.append(dict(type="disk", to_world=np.asarray(to_world, np.float64),
                                bsdf=bsdf if (bsdf is not None or emitter is not None) else dict(type="diffuse"), emitter=emitter,
                                flip_normals=flip_normals))
        return len(self.shapes) - 1

    def add_medium(self, sigma_a, sigma_s, phase="isotropic", g=0.0, strategy="balance"):
        """Homogeneous participating medium (src/medium/homogeneous.cpp); returns its index.  A shape refers to it through
        set_media(shape, interior=, exterior=)."""
        self.media.append(dict(sigma_a=tuple(float(x) for x in sigma_a), sigma_s=tuple(float(x) for x in sigma_s), phase=phase,
                               g=float(g), strategy=strategy))
        return len(self.media) - 1

    def set_camera_medium(self, medium):
        self.camera_medium = medium

    def set_media(self, shape, interior=None, exterior=None):
        self.shapes[shape]["interior"], self.shapes[shape]["exterior"] = interior, exterior

    def add_point_light(self, pos, intensity):
        self.emitters.append(dict(type="point", pos=tuple(pos), intensity=tuple(intensity)))

    def add_constant_env(self, radiance):
        self.emitters.append(dict(type="constant", radiance=tuple(radiance)))

    def add_envmap(self, image, scale=1.0, to_world=None, pyramid=None):
        """`envmap` emitter (src/emitters/envmap.cpp) from a latitude-longitude float RGB image [H, W, 3].  `pyramid` = the MIP
        levels the reference builds from it (list of [h, w, 3] arrays: half-precision values, Lanczos-downsampled,
        envmap.cpp:152-178) — hosts of the C ABI pass them on; the Mitsuba plugin builds them with the reference's own MIPMap."""
        self.emitters.append(dict(type="envmap", image=np.asarray(image, np.float32), scale=float(scale),
                                  to_world=np.eye(4) if to_world is None else np.asarray(to_world, np.float64), pyramid=pyramid))

    def set_sensor(self, to_world, fov, width, height, rfilter="box", spp=4, seed=0, sampler="independent", fov_axis="x",
                   aperture_radius=0.0, focus_distance=1.0):
        """aperture_radius > 0 selects the `thinlens` sensor (src/sensors/thinlens.cpp), else `perspective`."""
        self.sensor = dict(to_world=np.asarray(to_world, np.float64), fov=fov, width=width, height=height, rfilter=rfilter,
                           fov_axis=fov_axis, aperture_radius=float(aperture_radius), focus_distance=float(focus_distance))
        self.sampler = dict(kind=sampler, spp=spp, seed=seed)

    def set_integrator(self, kind="path", **kw):
        self.integrator = dict(kind=kind, **kw)

    # -- GPU back end ----------------------------------------------------------------------------
    @staticmethod
    def _bsdf_params(b):
        """dict -> (type id, parameter vector) in the layout include/mtsb200.h documents."""
        t = b["type"]
        two = B.BSDF_TWOSIDED_FLAG if b.get("twosided") else 0
        if t == "diffuse":
            return B.BSDF_DIFFUSE | two, list(b.get("reflectance", (0.5, 0.5, 0.5)))
        if t == "null":
            return B.BSDF_NULL, [0, 0, 0]
        if t == "passthrough":          # src/bsdfs/null.cpp: the boundary of an index-matched medium
            return B.BSDF_PASSTHROUGH, [0, 0, 0]
        if t == "dielectric":
            eta = b.get("int_ior", 1.5046) / b.get("ext_ior", 1.000277)
            return B.BSDF_DIELECTRIC, [eta] + list(b.get("specular_reflectance", (1, 1, 1))) + list(b.get("specular_transmittance", (1, 1, 1)))
        if t == "plastic":
            eta = b.get("int_ior", 1.49) / b.get("ext_ior", 1.000277)
            from .hostmath import fresnel_diffuse_reflectance
            fdr_int = fresnel_diffuse_reflectance(1.0 / eta)
            fdr_ext = fresnel_diffuse_reflectance(eta)
            dr = list(b.get("diffuse_reflectance", (0.5, 0.5, 0.5)))
            if b.get("texture") is not None and b["texture"].get("pyramid") is not None:
                # plastic.cpp:202-205 weighs the lobes by Texture::getAverage(): the mean of the texture's full-resolution level
                dr = [float(x) for x in np.asarray(b["texture"]["pyramid"][0], np.float64).reshape(-1, 3).mean(axis=0)]
            sr = list(b.get("specular_reflectance", (1, 1, 1)))
            lum = lambda c: 0.212671 * c[0] + 0.715160 * c[1] + 0.072169 * c[2]   # Spectrum::getLuminance (RGB build)
            d_avg, s_avg = lum(dr), lum(sr)
            ssw = s_avg / (d_avg + s_avg)
            return B.BSDF_PLASTIC | two, [eta] + dr + sr + [fdr_int, fdr_ext, ssw, 1.0 if b.get("nonlinear") else 0.0]
        if t == "roughconductor":
            a = b.get("alpha", 0.1)
            au, av = b.get("alpha_u", a), b.get("alpha_v", a)
            eta = list(b.get("eta", (0.2004, 0.9240, 1.1022)))     # Cu @ RGB, see hostmath.CU_ETA_K
            k = list(b.get("k", (3.9129, 2.4528, 2.1421)))
            ext = b.get("ext_eta", 1.000277)
            eta = [e / ext for e in eta]
            k = [e / ext for e in k]
            dist = dict(beckmann=0, ggx=1)[b.get("distribution", "beckmann")]
            return B.BSDF_ROUGHCONDUCTOR | two, [au, av] + eta + k + list(b.get("specular_reflectance", (1, 1, 1))) + [dist, 1.0 if b.get("sample_visible", True) else 0.0]
        if t == "conductor":
            ext = b.get("ext_eta", 1.000277)
            eta = [e / ext for e in b.get("eta", (0.2004, 0.9240, 1.1022))]
            k = [e / ext for e in b.get("k", (3.9129, 2.4528, 2.1421))]
            return B.BSDF_CONDUCTOR | two, eta + k + list(b.get("specular_reflectance", (1, 1, 1)))
        if t == "roughdielectric":
            a = b.get("alpha", 0.1)
            au, av = b.get("alpha_u", a), b.get("alpha_v", a)
            eta = b.get("int_ior", 1.5046) / b.get("ext_ior", 1.000277)
            dist = dict(beckmann=0, ggx=1)[b.get("distribution", "beckmann")]
            return B.BSDF_ROUGHDIELECTRIC, ([au, av, eta] + list(b.get("specular_reflectance", (1, 1, 1))) + list(b.get("specular_transmittance", (1, 1, 1)))
                                            + [0, 0, dist, 1.0 if b.get("sample_visible", True) else 0.0])
        if t == "roughplastic":
            # roughplastic.cpp:199-303; the transmittance slices come from the renderer's data/microfacet tables (MICROFACET_DIR)
            from . import hostmath as H
            eta = np.float32(b.get("int_ior", 1.49)) / np.float32(b.get("ext_ior", 1.000277))
            alpha = b.get("alpha", 0.1)
            dist = b.get("distribution", "beckmann")
            dr = list(b.get("diffuse_reflectance", (0.5, 0.5, 0.5)))
            sr = list(b.get("specular_reflectance", (1, 1, 1)))
            lum = lambda c: 0.212671 * c[0] + 0.715160 * c[1] + 0.072169 * c[2]
            ssw = lum(sr) / (lum(dr) + lum(sr))
            if MICROFACET_DIR is None:
                raise ValueError("roughplastic needs scenes.MICROFACET_DIR (the renderer's data/microfacet directory)")
            table, fdr = H.rough_transmittance_slices(H.load_rough_transmittance(os.path.join(MICROFACET_DIR, dist + ".dat")), eta, alpha)
            b["_table"] = table
            return B.BSDF_ROUGHPLASTIC | two, [alpha, float(eta)] + dr + sr + [ssw, fdr, 1.0 if b.get("nonlinear") else 0.0,
                                                                        dict(beckmann=0, ggx=1)[dist], 1.0 if b.get("sample_visible", True) else 0.0]
        raise ValueError("BSDF outside the GPU subset: %s" % t)

    def upload(self, ctx, builder=B.BUILDER_LBVH, ply_images=None):
        """ply_images: {shape index: binary PLY file image} — those mesh shapes are decoded on the device (Scene.add_ply) instead of
        being passed as arrays."""
        """Creates and commits a binding.Scene.  Emitter order = Mitsuba's: Scene::addChild registers a shape's area
        emitter when the shape is added, free emitters in XML order — free emitters are emitted first in to_xml()."""
        sc = B.Scene(ctx)
        for e in self.emitters:
            if e["type"] == "point":
                sc.add_emitter(B.EMITTER_POINT, radiance=e["intensity"], pos=e["pos"])
            elif e["type"] == "envmap":
                if e["pyramid"] is None:
                    raise ValueError("envmap: pass the MIP pyramid (SceneDesc.add_envmap(..., pyramid=...))")
                eid = sc.add_emitter(B.EMITTER_ENVMAP, radiance=(0, 0, 0))
                sc.set_envmap(eid, e["pyramid"], e["scale"], e["to_world"])
            else:
                sc.add_emitter(B.EMITTER_CONSTANT, radiance=e["radiance"])
        for m in self.media:
            # `single` samples with sigma_t of the lowest-variance channel, the smallest one (homogeneous.cpp:190-204)
            density = float(min(np.float32(a) + np.float32(b) for a, b in zip(m["sigma_a"], m["sigma_s"]))) if m["strategy"] == "single" else m.get("sampling_density", 0.0)
            sc.add_medium(m["sigma_a"], m["sigma_s"], strategy=dict(balance=0, single=1, manual=1)[m["strategy"]], sampling_density=density,
                          phase=dict(isotropic=0, hg=1)[m["phase"]], g=m["g"])
        for k, s in enumerate(self.shapes):
            bd = s["bsdf"]
            if s["emitter"] is not None and bd is None:
                bd = dict(type="null")
            tid, params = self._bsdf_params(bd)
            bid = sc.add_bsdf(tid, params)
            if bd.get("_table") is not None:
                sc.set_bsdf_table(bid, bd["_table"])
            if bd.get("texture") is not None:
                t = bd["texture"]
                if t.get("pyramid") is None:
                    raise ValueError("bitmap texture: pass the MIP pyramid (texture=dict(image=..., pyramid=[...]))")
                tid_ = sc.add_texture(t["pyramid"], wrap_u=t.get("wrap_u", t.get("wrap", "repeat")), wrap_v=t.get("wrap_v", t.get("wrap", "repeat")),
                                      filter_type=t.get("filter", "ewa"), max_anisotropy=t.get("max_anisotropy", 20.0),
                                      uv_scale=t.get("uv_scale", (1, 1)), uv_offset=t.get("uv_offset", (0, 0)))
                sc.set_bsdf_texture(bid, tid_)
            eid = -1
            if s["emitter"] is not None:
                eid = sc.add_emitter(B.EMITTER_AREA, radiance=s["emitter"])
            if s["type"] == "mesh" and ply_images is not None and k in ply_images:
                if eid >= 0:
                    raise ValueError("a device-decoded PLY mesh cannot carry an area emitter")
                sc.add_ply(ply_images[k], bsdf=bid)
            elif s["type"] == "mesh":
                sc.add_mesh(s["V"], s["F"], normals=s["N"], texcoords=s.get("UV"), bsdf=bid, emitter=eid)
            elif s["type"] == "sphere":
                sc.add_sphere(translate(*s["center"]), s["radius"], flip_normals=s["flip_normals"], bsdf=bid, emitter=eid)
            elif s["type"] == "disk":
                sc.add_disk(s["to_world"], flip_normals=s["flip_normals"], bsdf=bid, emitter=eid)
            else:
                sc.add_rectangle(s["to_world"], flip_normals=s["flip_normals"], bsdf=bid, emitter=eid)
            if s.get("interior") is not None or s.get("exterior") is not None:
                sc.set_shape_media(k, -1 if s.get("interior") is None else s["interior"],
                                   -1 if s.get("exterior") is None else s["exterior"])
        if self.sensor is not None and self.sensor.get("aperture_radius", 0.0) > 0:
            # the thinlens sensor's special shape: a disk of the aperture's radius in the sensor's frame (thinlens.cpp:522-534)
            sc.add_special_disk(np.asarray(self.sensor["to_world"], np.float64) @ scale(self.sensor["aperture_radius"]))
        if self.sensor is not None:
            # Scene::getAABB() includes the sensor's AABB (the camera position) when the environment emitter is sized
            M = np.asarray(self.sensor["to_world"], np.float64)
            r = float(self.sensor.get("aperture_radius", 0.0))
            # perspective: the camera position; thinlens: the box of its aperture (thinlens.cpp:515-519, getSpatialBounds)
            corners = np.array([[sx * r, sy * r, 0.0, 1.0] for sx in (-1, 1) for sy in (-1, 1)]) @ M.T
            sc.expand_bounds(corners[:, 0:3].min(axis=0).astype(np.float32), corners[:, 0:3].max(axis=0).astype(np.float32))
        sc.commit(builder)
        return sc

    def camera(self):
        s = self.sensor
        cam = make_camera(s["to_world"], s["fov"], s["width"], s["height"], fov_axis=s["fov_axis"])
        cam.aperture_radius, cam.focus_distance = s.get("aperture_radius", 0.0), s.get("focus_distance", 1.0)
        return cam

    def film(self, channels=5):
        s = self.sensor
        return make_film(s["width"], s["height"], s["rfilter"], channels=channels)

    def integrator_desc(self):
        return make_integrator(**self.integrator)

    def render(self, scene, seed=None, channels=5):
        """binding.Render for this description on an uploaded scene (sensor medium applied)."""
        rd = B.Render(scene, self.camera(), self.film(channels=channels), self.integrator_desc(), seed=self.sampler["seed"] if seed is None else seed)
        if self.camera_medium is not None:
            rd.set_camera_medium(self.camera_medium)
        if self.sampler["kind"] == "sobol":
            if SOBOL_TABLES is None:
                raise ValueError("sobol sampler: set scenes.SOBOL_TABLES (the reference's direction numbers, binding.Render.set_sobol)")
            rd.set_sobol(SOBOL_TABLES)
        elif self.sampler["kind"] in ("ldsampler", "ldctr"):
            # "ldctr": the oracle's counter sampler in its low-discrepancy mode (oracle/ctrsampler.cpp) — the same numbers, draw for draw
            rd.set_ldsampler(self.sampler.get("dimension", 4), self.sampler["spp"])
        return rd

    # -- reference back end ----------------------------------------------------------------------
    @staticmethod
    def _bsdf_xml(b):
        t = b["type"]
        rgb = lambda name, v: '<rgb name="%s" value="%s"/>' % (name, ", ".join(repr(float(x)) for x in v))
        if t == "null":
            return ""
        if t == "passthrough":          # index-matched medium boundary (src/bsdfs/null.cpp)
            return '<bsdf type="null"/>'
        inner = ""

        def tex_xml(name, x):
            # `bitmap` texture (src/textures/bitmap.cpp); to_xml() has written the image and set its file name
            return ('<texture type="bitmap" name="%s"><string name="filename" value="%s"/>'
                    '<string name="wrapModeU" value="%s"/><string name="wrapModeV" value="%s"/><string name="filterType" value="%s"/>'
                    '<float name="maxAnisotropy" value="%r"/><float name="uscale" value="%r"/><float name="vscale" value="%r"/>'
                    '<float name="uoffset" value="%r"/><float name="voffset" value="%r"/><boolean name="cache" value="false"/></texture>') % (
                name, x["filename"], x.get("wrap_u", x.get("wrap", "repeat")), x.get("wrap_v", x.get("wrap", "repeat")), x.get("filter", "ewa"),
                float(x.get("max_anisotropy", 20.0)), float(x.get("uv_scale", (1, 1))[0]), float(x.get("uv_scale", (1, 1))[1]),
                float(x.get("uv_offset", (0, 0))[0]), float(x.get("uv_offset", (0, 0))[1]))
        if t == "diffuse" and b.get("texture") is not None:
            inner = '<bsdf type="diffuse">%s</bsdf>' % tex_xml("reflectance", b["texture"])
        elif t == "plastic" and b.get("texture") is not None:
            inner = '<bsdf type="plastic"><float name="intIOR" value="%r"/><float name="extIOR" value="%r"/>%s%s<boolean name="nonlinear" value="%s"/></bsdf>' % (
                b.get("int_ior", 1.49), b.get("ext_ior", 1.000277), tex_xml("diffuseReflectance", b["texture"]),
                rgb("specularReflectance", b.get("specular_reflectance", (1, 1, 1))), "true" if b.get("nonlinear") else "false")
        elif t == "diffuse":
            inner = '<bsdf type="diffuse">%s</bsdf>' % rgb("reflectance", b.get("reflectance", (0.5, 0.5, 0.5)))
        elif t == "dielectric":
            inner = '<bsdf type="dielectric"><float name="intIOR" value="%r"/><float name="extIOR" value="%r"/></bsdf>' % (
                b.get("int_ior", 1.5046), b.get("ext_ior", 1.000277))
        elif t == "plastic":
            inner = '<bsdf type="plastic"><float name="intIOR" value="%r"/><float name="extIOR" value="%r"/>%s%s<boolean name="nonlinear" value="%s"/></bsdf>' % (
                b.get("int_ior", 1.49), b.get("ext_ior", 1.000277), rgb("diffuseReflectance", b.get("diffuse_reflectance", (0.5, 0.5, 0.5))),
                rgb("specularReflectance", b.get("specular_reflectance", (1, 1, 1))), "true" if b.get("nonlinear") else "false")
        elif t == "roughplastic":
            inner = ('<bsdf type="roughplastic"><string name="distribution" value="%s"/><float name="alpha" value="%r"/><float name="intIOR" value="%r"/>'
                     '<float name="extIOR" value="%r"/>%s%s<boolean name="nonlinear" value="%s"/><boolean name="sampleVisible" value="%s"/></bsdf>') % (
                b.get("distribution", "beckmann"), b.get("alpha", 0.1), b.get("int_ior", 1.49), b.get("ext_ior", 1.000277),
                rgb("diffuseReflectance", b.get("diffuse_reflectance", (0.5, 0.5, 0.5))), rgb("specularReflectance", b.get("specular_reflectance", (1, 1, 1))),
                "true" if b.get("nonlinear") else "false", "true" if b.get("sample_visible", True) else "false")
        elif t == "roughconductor":
            a = b.get("alpha", 0.1)
            inner = ('<bsdf type="roughconductor"><string name="distribution" value="%s"/><float name="alphaU" value="%r"/>'
                     '<float name="alphaV" value="%r"/>%s%s<float name="extEta" value="%r"/><boolean name="sampleVisible" value="%s"/></bsdf>') % (
                b.get("distribution", "beckmann"), b.get("alpha_u", a), b.get("alpha_v", a),
                rgb("eta", b.get("eta", (0.2004, 0.9240, 1.1022))), rgb("k", b.get("k", (3.9129, 2.4528, 2.1421))),
                b.get("ext_eta", 1.000277), "true" if b.get("sample_visible", True) else "false")
        elif t == "conductor":
            inner = '<bsdf type="conductor">%s%s%s<float name="extEta" value="%r"/></bsdf>' % (
                rgb("eta", b.get("eta", (0.2004, 0.9240, 1.1022))), rgb("k", b.get("k", (3.9129, 2.4528, 2.1421))),
                rgb("specularReflectance", b.get("specular_reflectance", (1, 1, 1))), b.get("ext_eta", 1.000277))
        elif t == "roughdielectric":
            a = b.get("alpha", 0.1)
            inner = ('<bsdf type="roughdielectric"><string name="distribution" value="%s"/><float name="alphaU" value="%r"/>'
                     '<float name="alphaV" value="%r"/><float name="intIOR" value="%r"/><float name="extIOR" value="%r"/>%s%s'
                     '<boolean name="sampleVisible" value="%s"/></bsdf>') % (
                b.get("distribution", "beckmann"), b.get("alpha_u", a), b.get("alpha_v", a), b.get("int_ior", 1.5046), b.get("ext_ior", 1.000277),
                rgb("specularReflectance", b.get("specular_reflectance", (1, 1, 1))), rgb("specularTransmittance", b.get("specular_transmittance", (1, 1, 1))),
                "true" if b.get("sample_visible", True) else "false")
        else:
            raise ValueError(t)
        if b.get("twosided"):
            inner = '<bsdf type="twosided">%s</bsdf>' % inner
        return inner

    def to_xml(self, directory, name="scene", film_format="pfm", mesh_format="ply"):
        """Writes <directory>/<name>.xml (+ one PLY per mesh, or with mesh_format="serialized" ONE `serialized` file holding every mesh,
        addressed by shapeIndex) and returns the XML path."""
        ser_meshes = []
        os.makedirs(directory, exist_ok=True)
        out = ['<?xml version="1.0" encoding="utf-8"?>', '<scene version="0.5.0">']
        ig = self.integrator
        out.append('<integrator type="%s">' % ig["kind"])
        for k, xmlk in (("max_depth", "maxDepth"), ("rr_depth", "rrDepth"), ("emitter_samples", "emitterSamples"), ("bsdf_samples", "bsdfSamples")):
            if k in ig:
                out.append('<integer name="%s" value="%d"/>' % (xmlk, ig[k]))
        for k, xmlk in (("strict_normals", "strictNormals"), ("hide_emitters", "hideEmitters")):
            if k in ig:
                out.append('<boolean name="%s" value="%s"/>' % (xmlk, "true" if ig[k] else "false"))
        out.append('</integrator>')
        for k, m in enumerate(self.media):
            ph = '<phase type="hg"><float name="g" value="%r"/></phase>' % m["g"] if m["phase"] == "hg" else '<phase type="isotropic"/>'
            out.append('<medium type="homogeneous" id="medium%d"><rgb name="sigmaA" value="%s"/><rgb name="sigmaS" value="%s"/>'
                       '<string name="strategy" value="%s"/>%s</medium>' % (k, ", ".join(repr(x) for x in m["sigma_a"]),
                                                                           ", ".join(repr(x) for x in m["sigma_s"]), m["strategy"], ph))
        if self.sensor is not None:
            s, sm = self.sensor, self.sampler
            if s.get("aperture_radius", 0.0) > 0:
                out.append('<sensor type="thinlens"><float name="apertureRadius" value="%r"/><float name="focusDistance" value="%r"/>'
                           '<float name="fov" value="%r"/><string name="fovAxis" value="%s"/>' % (s["aperture_radius"], s["focus_distance"], s["fov"], s["fov_axis"]))
            else:
                out.append('<sensor type="perspective"><float name="fov" value="%r"/><string name="fovAxis" value="%s"/>' % (s["fov"], s["fov_axis"]))
            out.append('<transform name="toWorld">%s</transform>' % _mat_xml(s["to_world"]))
            if sm["kind"] == "sobol":
                out.append('<sampler type="sobol"><integer name="sampleCount" value="%d"/></sampler>' % sm["spp"])
            elif sm["kind"] == "ldsampler":
                out.append('<sampler type="ldsampler"><integer name="sampleCount" value="%d"/><integer name="dimension" value="%d"/></sampler>' % (sm["spp"], sm.get("dimension", 4)))
            elif sm["kind"] == "ldctr":
                out.append('<sampler type="ctr"><integer name="sampleCount" value="%d"/><integer name="seed" value="%d"/><integer name="ldDimension" value="%d"/></sampler>' % (
                    sm["spp"], sm["seed"], sm.get("dimension", 4)))
            elif sm["kind"] == "ctr":
                out.append('<sampler type="ctr"><integer name="sampleCount" value="%d"/><integer name="seed" value="%d"/></sampler>' % (sm["spp"], sm["seed"]))
            else:
                out.append('<sampler type="independent"><integer name="sampleCount" value="%d"/></sampler>' % sm["spp"])
            out.append('<film type="hdrfilm"><integer name="width" value="%d"/><integer name="height" value="%d"/>'
                       '<rfilter type="%s"/><string name="fileFormat" value="%s"/><string name="pixelFormat" value="rgb"/>'
                       '<boolean name="banner" value="false"/></film>' % (s["width"], s["height"], s["rfilter"], film_format))
            if self.camera_medium is not None:
                out.append('<ref id="medium%d"/>' % self.camera_medium)
            out.append('</sensor>')
        for e in self.emitters:
            if e["type"] == "point":
                out.append('<emitter type="point"><point name="position" x="%r" y="%r" z="%r"/><rgb name="intensity" value="%s"/></emitter>' % (
                    e["pos"][0], e["pos"][1], e["pos"][2], ", ".join(repr(float(x)) for x in e["intensity"])))
            elif e["type"] == "envmap":
                fn = "%s_envmap.pfm" % name
                write_pfm(os.path.join(directory, fn), e["image"])
                out.append('<emitter type="envmap"><string name="filename" value="%s"/><float name="scale" value="%r"/><boolean name="cache" value="false"/>'
                           '<transform name="toWorld">%s</transform></emitter>' % (fn, e["scale"], _mat_xml(e["to_world"])))
            else:
                out.append('<emitter type="constant"><rgb name="radiance" value="%s"/></emitter>' % ", ".join(repr(float(x)) for x in e["radiance"]))
        for i, s in enumerate(self.shapes):
            if s["bsdf"] is not None and s["bsdf"].get("texture") is not None:
                s["bsdf"]["texture"]["filename"] = "%s_tex%d.pfm" % (name, i)
                write_pfm(os.path.join(directory, s["bsdf"]["texture"]["filename"]), s["bsdf"]["texture"]["image"])
            em = ""
            if s["emitter"] is not None:
                em = '<emitter type="area"><rgb name="radiance" value="%s"/></emitter>' % ", ".join(repr(float(x)) for x in s["emitter"])
            bs = self._bsdf_xml(s["bsdf"]) if s["bsdf"] is not None else ""
            for side in ("interior", "exterior"):
                if s.get(side) is not None:
                    bs += '<ref name="%s" id="medium%d"/>' % (side, s[side])
            if s["type"] == "mesh" and mesh_format == "serialized":
                face = "" if s["N"] is not None else '<boolean name="faceNormals" value="true"/>'
                out.append('<shape type="serialized"><string name="filename" value="%s.serialized"/><integer name="shapeIndex" value="%d"/>%s%s%s</shape>' % (
                    name, len(ser_meshes), face, bs, em))
                ser_meshes.append(dict(V=s["V"], F=s["F"], N=s["N"], UV=s.get("UV"), name="mesh%d" % i))
            elif s["type"] == "mesh":
                fn = "%s_mesh%d.ply" % (name, i)
                write_ply(os.path.join(directory, fn), s["V"], s["N"], s["F"], s.get("UV"))
                face = "" if s["N"] is not None else '<boolean name="faceNormals" value="true"/>'
                out.append('<shape type="ply"><string name="filename" value="%s"/>%s%s%s</shape>' % (fn, face, bs, em))
            elif s["type"] == "sphere":
                flip = '<boolean name="flipNormals" value="true"/>' if s["flip_normals"] else ""
                out.append('<shape type="sphere"><point name="center" x="%r" y="%r" z="%r"/><float name="radius" value="%r"/>%s%s%s</shape>' % (
                    s["center"][0], s["center"][1], s["center"][2], s["radius"], flip, bs, em))
            elif s["type"] == "disk":
                flip = '<boolean name="flipNormals" value="true"/>' if s["flip_normals"] else ""
                out.append('<shape type="disk"><transform name="toWorld">%s</transform>%s%s%s</shape>' % (_mat_xml(s["to_world"]), flip, bs, em))
            else:
                flip = '<boolean name="flipNormals" value="true"/>' if s["flip_normals"] else ""
                out.append('<shape type="rectangle"><transform name="toWorld">%s</transform>%s%s%s</shape>' % (_mat_xml(s["to_world"]), flip, bs, em))
        out.append('</scene>')
        if ser_meshes:
            with open(os.path.join(directory, name + ".serialized"), "wb") as f:
                f.write(serialized_image(ser_meshes))
        path = os.path.join(directory, name + ".xml")
        with open(path, "w") as f:
            f.write("\n".join(out))
        return path


# ------------------------------------------------------------------------------------------------ the configs
def cube_mesh(to_world):
    """A 12-triangle, 24-vertex box [-1,1]^3 with per-face normals (used instead of the `cube` plugin when the scene
    is driven through the C ABI; the Mitsuba plugin path reads the real `cube` TriMesh)."""
    V, N, F = [], [], []
    for axis in range(3):
        for sgn in (-1.0, 1.0):
            n = np.zeros(3)
            n[axis] = sgn
            u = np.zeros(3)
            u[(axis + 1) % 3] = 1.0
            v = np.cross(n, u)
            base = len(V)
            for a, b in ((-1, -1), (1, -1), (1, 1), (-1, 1)):
                V.append(n + a * u + b * v)
                N.append(n)
            F += [[base, base + 1, base + 2], [base + 2, base + 3, base]]
    V = np.asarray(V)
    N = np.asarray(N)
    M = np.asarray(to_world, dtype=np.float64)
    Vw = V @ M[:3, :3].T + M[:3, 3]
    Nw = N @ np.linalg.inv(M[:3, :3])
    Nw /= np.linalg.norm(Nw, axis=1, keepdims=True)
    return Vw.astype(np.float32), Nw.astype(np.float32), np.asarray(F, np.uint32)


def cornell_box(width=512, height=512, spp=64, max_depth=5, rfilter="box", sampler="independent", seed=0):
    """Config 1: 5 diffuse rectangles + a cube + a rectangular area light."""
    sd = SceneDesc()
    white = dict(type="diffuse", reflectance=(0.5, 0.5, 0.5))
    sd.add_rectangle(translate(0, 0, 0) @ rotate((1, 0, 0), -90), bsdf=white)
    sd.add_rectangle(translate(0, 2, 0) @ rotate((1, 0, 0), 90), bsdf=white)
    sd.add_rectangle(translate(0, 1, -1), bsdf=white)
    sd.add_rectangle(translate(-1, 1, 0) @ rotate((0, 1, 0), 90), bsdf=dict(type="diffuse", reflectance=(0.63, 0.065, 0.05)))
    sd.add_rectangle(translate(1, 1, 0) @ rotate((0, 1, 0), -90), bsdf=dict(type="diffuse", reflectance=(0.14, 0.45, 0.091)))
    V, N, F = cube_mesh(translate(0.3, 0.3, 0.2) @ rotate((0, 1, 0), 20) @ scale(0.3))
    sd.add_mesh(V, N, F, bsdf=white)
    sd.add_rectangle(translate(0, 1.98, 0) @ rotate((1, 0, 0), 90) @ scale(0.25), bsdf=None, emitter=(17, 12, 4))
    sd.set_sensor(look_at((0, 1, 3.9), (0, 1, 0), (0, 1, 0)), 39.3077, width, height, rfilter=rfilter, spp=spp, sampler=sampler, seed=seed)
    sd.set_integrator("path", max_depth=max_depth, rr_depth=5)
    return sd


def sphere_field(subdiv=6, nx=16, ny=8, width=1920, height=1080, spp=256, max_depth=8, rfilter="gaussian",
                 sampler="independent", seed=0):
    """Config 3/4/5: nx*ny displaced icospheres (20*4**subdiv triangles each) alternating GGX rough copper and bk7
    glass on a diffuse ground quad, one area light quad + constant environment.  subdiv=6, 16x8 -> 10 485 762 triangles."""
    sd = SceneDesc()
    base_V, base_F = icosphere(subdiv)
    metal_V, metal_N, metal_F, glass_V, glass_N, glass_F = [], [], [], [], [], []
    mo = go = 0
    for j in range(ny):
        for i in range(nx):
            idx = j * nx + i
            V = base_V * (1.0 + 0.05 * fbm(base_V, idx + 1))[:, None]
            N = vertex_normals(V, base_F)
            V = V * 0.45 + np.array([i - (nx - 1) / 2.0, 0.5, j - (ny - 1) / 2.0])
            if (i + j) % 2 == 0:
                metal_V.append(V); metal_N.append(N); metal_F.append(base_F + mo); mo += V.shape[0]
            else:
                glass_V.append(V); glass_N.append(N); glass_F.append(base_F + go); go += V.shape[0]
    sd.add_mesh(np.concatenate(metal_V), np.concatenate(metal_N), np.concatenate(metal_F),
                bsdf=dict(type="roughconductor", distribution="ggx", alpha=0.1))
    if glass_V:
        sd.add_mesh(np.concatenate(glass_V), np.concatenate(glass_N), np.concatenate(glass_F), bsdf=dict(type="dielectric"))
    g = max(nx, ny) * 0.75
    gV = np.array([[-g, 0, -g], [g, 0, -g], [g, 0, g], [-g, 0, g]], np.float32)
    gN = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    sd.add_mesh(gV, gN, np.array([[0, 2, 1], [0, 3, 2]], np.uint32), bsdf=dict(type="diffuse", reflectance=(0.5, 0.5, 0.5)))
    sd.add_rectangle(translate(0, 6, 0) @ rotate((1, 0, 0), 90) @ scale(2.0), bsdf=None, emitter=(30, 30, 30))
    sd.add_constant_env((0.5, 0.5, 0.5))
    sd.set_sensor(look_at((0, 5.5, -ny * 1.1 - 2), (0, 0.3, 0), (0, 1, 0)), 45.0, width, height, rfilter=rfilter, spp=spp,
                  sampler=sampler, seed=seed)
    sd.set_integrator("path", max_depth=max_depth, rr_depth=5)
    return sd


# ------------------------------------------------------------------------------------------------ ray batches (config 2)
def _uniform_sphere(u):
    z = 1.0 - 2.0 * u[:, 1]
    r = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    phi = 2.0 * math.pi * u[:, 0]
    return np.stack([r * np.cos(phi), r * np.sin(phi), z], axis=1)


def chord_rays(n, center=(0, 0, 0), radius=1.1, seed=42, mint=0.0):
    """Incoherent rays between two uniform points on a bounding sphere — the distribution of `mtsutil kdbench`
    (src/utils/kdbench.cpp:226-230).  mint=EPSILON reproduces kdbench's Ray(o,d,time) constructor default."""
    rng = np.random.Generator(np.random.Philox(seed))
    u = rng.random((n, 4), dtype=np.float32).astype(np.float64)
    c = np.asarray(center, dtype=np.float64)
    p1 = c + radius * _uniform_sphere(u[:, 0:2])
    p2 = c + radius * _uniform_sphere(u[:, 2:4])
    d = p2 - p1
    d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-20)
    rays = np.zeros(n, B.ray_dtype)
    rays["o"], rays["d"] = p1.astype(np.float32), d.astype(np.float32)
    rays["mint"], rays["maxt"] = np.float32(mint), np.float32(np.inf)
    return rays


def pinhole_rays(res, origin=(0, 0, -3.0), target=(0, 0, 0), fov_deg=45.0):
    """Coherent primary rays of a res x res pinhole camera, row-major."""
    M = look_at(origin, target, (0, 1, 0))
    t = math.tan(math.radians(fov_deg) / 2.0)
    xs = (np.arange(res) + 0.5) / res * 2.0 - 1.0
    px, py = np.meshgrid(xs, -xs)
    d = np.stack([px.ravel() * t, py.ravel() * t, np.ones(res * res)], axis=1)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d @ M[:3, :3].T
    rays = np.zeros(res * res, B.ray_dtype)
    rays["o"] = np.asarray(origin, np.float32)
    rays["d"] = d.astype(np.float32)
    rays["mint"], rays["maxt"] = np.float32(EPSILON), np.float32(np.inf)
    return rays


def shadow_rays(points, light_center=(0, 3.0, 0), light_half=0.2, seed=43):
    """Shadow segments from surface points to uniform points on a light quad, with the reference's conventions:
    mint = Epsilon (scaled adaptively on the device) and maxt = dist * (1 - ShadowEpsilon)  (src/librender/scene.cpp:871-873)."""
    n = points.shape[0]
    rng = np.random.Generator(np.random.Philox(seed))
    u = rng.random((n, 2), dtype=np.float32).astype(np.float64)
    lc = np.asarray(light_center, dtype=np.float64)
    lp = lc + np.stack([(u[:, 0] * 2 - 1) * light_half, np.zeros(n), (u[:, 1] * 2 - 1) * light_half], axis=1)
    d = lp - points.astype(np.float64)
    dist = np.linalg.norm(d, axis=1)
    d /= np.maximum(dist[:, None], 1e-20)
    rays = np.zeros(n, B.ray_dtype)
    rays["o"], rays["d"] = points.astype(np.float32), d.astype(np.float32)
    rays["mint"] = np.float32(EPSILON)
    rays["maxt"] = (dist.astype(np.float32) * np.float32(1 - SHADOW_EPSILON)).astype(np.float32)
    return rays


# ------------------------------------------------------------------------------------------------ small test scenes
MATERIALS = [
    dict(type="diffuse", reflectance=(0.6, 0.3, 0.2)),
    dict(type="dielectric"),
    dict(type="plastic", diffuse_reflectance=(0.2, 0.3, 0.6)),
    dict(type="plastic", diffuse_reflectance=(0.7, 0.6, 0.1), nonlinear=True, int_ior=1.9),
    dict(type="roughconductor", distribution="ggx", alpha=0.1),
    dict(type="roughconductor", distribution="beckmann", alpha=0.3),
    dict(type="roughconductor", distribution="ggx", alpha_u=0.05, alpha_v=0.3, sample_visible=False),
    dict(type="diffuse", reflectance=(0.4, 0.7, 0.4), twosided=True),
    dict(type="conductor", eta=(0.143, 0.375, 1.442), k=(3.983, 2.386, 1.603), specular_reflectance=(0.9, 0.95, 1.0)),
    dict(type="roughdielectric", distribution="ggx", alpha=0.15, specular_transmittance=(0.9, 1.0, 0.9)),
    dict(type="roughdielectric", distribution="beckmann", alpha_u=0.1, alpha_v=0.25, int_ior=1.33, sample_visible=False),
]


# `roughplastic`: both distributions, nonlinear colour shift, sampling of all normals, two-sided
ROUGH_PLASTICS = [
    dict(type="roughplastic", distribution="beckmann", alpha=0.1, diffuse_reflectance=(0.2, 0.3, 0.6)),
    dict(type="roughplastic", distribution="ggx", alpha=0.3, diffuse_reflectance=(0.7, 0.6, 0.1), nonlinear=True, int_ior=1.9),
    dict(type="roughplastic", distribution="beckmann", alpha=0.25, sample_visible=False, specular_reflectance=(0.8, 0.9, 1.0)),
    dict(type="roughplastic", distribution="ggx", alpha=0.05, diffuse_reflectance=(0.5, 0.2, 0.2), twosided=True),
]


def analytic_scene(width=48, height=32, spp=4, max_depth=6, sampler="independent", seed=0):
    """Analytic `sphere` and `disk` shapes (src/shapes/sphere.cpp, disk.cpp) beside meshes: a glass, a diffuse and a GGX sphere,
    a tilted plastic disk and a flipped diffuse one, on a mesh floor under a mesh area light, a point light and an environment."""
    sd = SceneDesc()
    sd.add_point_light((-3.0, 4.0, -2.0), (20, 18, 15))
    sd.add_constant_env((0.3, 0.35, 0.5))
    sd.add_sphere((-1.1, 0.5, 0.2), 0.5, bsdf=dict(type="dielectric"))
    sd.add_sphere((0.1, 0.4, -0.3), 0.4, bsdf=dict(type="diffuse", reflectance=(0.6, 0.3, 0.2)))
    sd.add_sphere((1.2, 0.45, 0.4), 0.45, bsdf=dict(type="roughconductor", distribution="ggx", alpha=0.1))
    sd.add_sphere((0.3, 1.5, 1.4), 0.35, bsdf=dict(type="diffuse", reflectance=(0.2, 0.5, 0.3)), flip_normals=True)
    sd.add_disk(translate(-0.4, 0.8, 1.2) @ rotate((1, 0, 0), -60) @ scale(0.6), bsdf=dict(type="plastic", diffuse_reflectance=(0.2, 0.3, 0.6)))
    sd.add_disk(translate(1.6, 0.9, -0.8) @ rotate((0, 1, 0), 70) @ scale(0.5), bsdf=dict(type="diffuse", twosided=True), flip_normals=True)
    g = 4.0
    gV = np.array([[-g, 0, -g], [g, 0, -g], [g, 0, g], [-g, 0, g]], np.float32)
    gN = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    sd.add_mesh(gV, gN, np.array([[0, 2, 1], [0, 3, 2]], np.uint32), bsdf=dict(type="diffuse", reflectance=(0.5, 0.5, 0.5)))
    lV, lN, lF = displaced_sphere(1, seed=99, radius=0.4, center=(0.5, 2.6, 0.3), amplitude=0.0)
    sd.add_mesh(lV, lN, lF, bsdf=None, emitter=(25, 22, 18))
    sd.set_sensor(look_at((0.3, 3.0, -4.5), (0, 0.4, 0), (0, 1, 0)), 40.0, width, height, rfilter="box", spp=spp, sampler=sampler, seed=seed)
    sd.set_integrator("path", max_depth=max_depth, rr_depth=3)
    return sd


def analytic_lights_scene(width=48, height=32, spp=4, max_depth=6, sampler="independent", seed=0):
    """Area lights on analytic shapes: a small emitting `sphere` above the scene (cone sampling from outside,
    sphere.cpp:286-355), a large emitting sphere the camera and the scene sit INSIDE (uniform-sphere branch, flipped normals so
    that it shines inwards), and an emitting `disk` (Disk::samplePosition); diffuse, glass and GGX receivers."""
    sd = SceneDesc()
    sd.add_sphere((-1.2, 2.2, 0.3), 0.25, emitter=(30, 26, 20))
    sd.add_sphere((0.0, 1.0, 0.0), 9.0, emitter=(0.15, 0.2, 0.3), flip_normals=True)
    sd.add_disk(translate(1.4, 1.8, 0.6) @ rotate((1, 0, 0), 110) @ scale(0.35), emitter=(18, 14, 9))
    sd.add_sphere((-1.1, 0.5, 0.2), 0.5, bsdf=dict(type="dielectric"))
    V, N, F = displaced_sphere(2, seed=5, radius=0.45, center=(0.2, 0.45, -0.2))
    sd.add_mesh(V, N, F, bsdf=dict(type="diffuse", reflectance=(0.6, 0.3, 0.2)))
    sd.add_sphere((1.3, 0.45, 0.5), 0.45, bsdf=dict(type="roughconductor", distribution="ggx", alpha=0.1))
    g = 4.0
    gV = np.array([[-g, 0, -g], [g, 0, -g], [g, 0, g], [-g, 0, g]], np.float32)
    gN = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    sd.add_mesh(gV, gN, np.array([[0, 2, 1], [0, 3, 2]], np.uint32), bsdf=dict(type="diffuse", reflectance=(0.5, 0.5, 0.5)))
    sd.set_sensor(look_at((0.3, 3.0, -4.5), (0, 0.4, 0), (0, 1, 0)), 40.0, width, height, rfilter="box", spp=spp, sampler=sampler, seed=seed)
    sd.set_integrator("path", max_depth=max_depth, rr_depth=3)
    return sd


def texture_image(size=64, seed=9):
    """A colour texture with fine detail (checker x stripes + noise, values in (0, 0.9]): minification makes the MIP levels matter."""
    y, x = np.meshgrid(np.arange(size), np.arange(size), indexing="ij")
    chk = ((x // 4 + y // 4) % 2).astype(np.float32)
    img = np.stack([0.15 + 0.7 * chk, 0.2 + 0.6 * ((x // 2) % 2), 0.25 + 0.5 * ((y // 8) % 2)], axis=-1).astype(np.float32)
    rng = np.random.default_rng(seed)
    return np.clip(img * (1.0 + 0.1 * rng.standard_normal((size, size, 1))), 0.02, 0.9).astype(np.float32)


def textured_scene(width=48, height=32, spp=4, max_depth=5, sampler="independent", seed=0, pyramids=None, kind="path", plastic=False):
    """`bitmap` textures on the reflectance of diffuse BSDFs: a floor mesh with texture coordinates that repeat the
    texture 6 x 6 times (EWA, seen at a grazing angle: anisotropic footprints), a wall `rectangle` (trilinear, clamp, uv scale and
    offset) and a displaced sphere with spherical coordinates (bilinear, mirror); pyramids = the three MIP pyramids.
    plastic=True: the wall and the sphere carry the texture on the diffuseReflectance of a `plastic` BSDF instead (the sphere's
    with nonlinear colour shifts)."""
    img = texture_image()
    pyramids = pyramids or [None, None, None]
    sd = SceneDesc()
    sd.add_point_light((-3.0, 4.0, -2.0), (20, 18, 15))
    sd.add_constant_env((0.3, 0.35, 0.5))
    g = 4.0
    gV = np.array([[-g, 0, -g], [g, 0, -g], [g, 0, g], [-g, 0, g]], np.float32)
    gN = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    gUV = np.array([[0, 0], [6, 0], [6, 6], [0, 6]], np.float32)
    sd.add_mesh(gV, gN, np.array([[0, 2, 1], [0, 3, 2]], np.uint32), UV=gUV,
                bsdf=dict(type="diffuse", texture=dict(image=img, pyramid=pyramids[0], wrap="repeat", filter="ewa")))
    mat = "plastic" if plastic else "diffuse"
    sd.add_rectangle(translate(0, 1.5, 2.5) @ rotate((0, 1, 0), 180) @ scale(2.5, 1.5, 1.0),
                     bsdf=dict(type=mat, texture=dict(image=img, pyramid=pyramids[1], wrap="clamp", filter="trilinear", uv_scale=(3, 2), uv_offset=(-0.4, 0.1))))
    c = np.array((0.3, 0.7, 0.2))
    V, N, F = displaced_sphere(3, seed=33, radius=0.7, center=tuple(c))
    dn = (V - c) / np.linalg.norm(V - c, axis=1, keepdims=True)
    UV = np.stack([np.arctan2(dn[:, 2], dn[:, 0]) / (2 * math.pi) + 0.5, np.arccos(np.clip(dn[:, 1], -1, 1)) / math.pi], axis=1) * 3.0
    sd.add_mesh(V, N, F, UV=UV, bsdf=dict(type=mat, nonlinear=plastic, int_ior=1.7, texture=dict(image=img, pyramid=pyramids[2], wrap="mirror", filter="bilinear")))
    V, N, F = displaced_sphere(2, seed=34, radius=0.4, center=(-1.3, 0.4, -0.4))
    sd.add_mesh(V, N, F, bsdf=dict(type="dielectric"))
    sd.set_sensor(look_at((0.3, 1.3, -4.5), (0, 0.7, 0), (0, 1, 0)), 50.0, width, height, rfilter="box", spp=spp, sampler=sampler, seed=seed)
    sd.set_integrator(kind, max_depth=max_depth, rr_depth=3)
    return sd


def envmap_scene(width=48, height=32, spp=4, max_depth=5, sampler="independent", seed=0, pyramid=None, kind="path"):
    """`envmap` emitter: the synthetic sky of sky_image() rotated about two axes and scaled, as the only light beside a
    small mesh light; diffuse, glass, GGX and plastic spheres on a floor; the camera sees the sky directly (filtered lookups of
    sensor rays, envmap.cpp:395-409) and through the glass ball."""
    sd = SceneDesc()
    sd.add_envmap(sky_image(), scale=1.3, to_world=rotate((0, 1, 0), 35) @ rotate((1, 0, 0), 20), pyramid=pyramid)
    mats = [dict(type="diffuse", reflectance=(0.6, 0.5, 0.4)), dict(type="dielectric"), dict(type="roughconductor", distribution="ggx", alpha=0.15),
            dict(type="plastic", diffuse_reflectance=(0.2, 0.3, 0.6))]
    for i, m in enumerate(mats):
        V, N, F = displaced_sphere(2, seed=70 + i, radius=0.45, center=(i - 1.5, 0.5, 0.2 * (i % 2)))
        sd.add_mesh(V, N, F, bsdf=m)
    g = 3.0
    gV = np.array([[-g, 0, -g], [g, 0, -g], [g, 0, g], [-g, 0, g]], np.float32)
    gN = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    sd.add_mesh(gV, gN, np.array([[0, 2, 1], [0, 3, 2]], np.uint32), bsdf=dict(type="diffuse", reflectance=(0.5, 0.5, 0.5)))
    lV, lN, lF = displaced_sphere(1, seed=99, radius=0.2, center=(0.5, 2.2, 0.3), amplitude=0.0)
    sd.add_mesh(lV, lN, lF, bsdf=None, emitter=(12, 10, 8))
    sd.set_sensor(look_at((0.3, 1.6, -4.5), (0, 0.9, 0), (0, 1, 0)), 50.0, width, height, rfilter="box", spp=spp, sampler=sampler, seed=seed)
    sd.set_integrator(kind, max_depth=max_depth, rr_depth=3)
    return sd


def fog_scene(width=48, height=32, spp=4, max_depth=6, sampler="independent", seed=0, kind="volpath_simple", haze=False, sphere_light=False, envmap=False, envmap_pyramid=None,
              ball_bsdf=None, aperture_radius=0.0):
    """Participating media: a ball of isotropic fog behind an index-matched boundary (no BSDF: `null`), a
    glass ball filled with a forward-scattering (HG, g = 0.6) coloured medium, a diffuse ball, on a mesh floor under a mesh
    area light, a point light and a constant environment; camera outside the media.  haze=True wraps all of it in a large
    index-matched ball of thin haze (`single` sampling strategy) with the camera inside: paths start in a medium, the balls'
    boundaries separate two media, and rays leave the haze through its boundary to reach the environment."""
    sd = SceneDesc()
    sd.add_point_light((-3.0, 4.0, -2.0), (20, 18, 15))
    if envmap:      # the synthetic sky of envmap_scene instead of the constant environment (seen through the media)
        sd.add_envmap(sky_image(), scale=1.3, to_world=rotate((0, 1, 0), 35) @ rotate((1, 0, 0), 20), pyramid=envmap_pyramid)
    else:
        sd.add_constant_env((0.3, 0.35, 0.5))
    fog = sd.add_medium((0.2, 0.2, 0.2), (1.6, 1.8, 2.2))
    milk = sd.add_medium((0.05, 0.4, 0.8), (3.0, 2.5, 1.5), phase="hg", g=0.6)
    V, N, F = displaced_sphere(2, seed=41, radius=0.6, center=(-1.0, 0.65, 0.1), amplitude=0.0)
    a = sd.add_mesh(V, N, F, bsdf=dict(type="passthrough"))
    hz = sd.add_medium((0.02, 0.02, 0.02), (0.08, 0.09, 0.1), strategy="single") if haze else None
    sd.set_media(a, interior=fog, exterior=hz)
    V, N, F = displaced_sphere(2, seed=42, radius=0.5, center=(0.4, 0.55, -0.2), amplitude=0.0)
    b = sd.add_mesh(V, N, F, bsdf=dict(type="dielectric"))
    sd.set_media(b, interior=milk, exterior=hz)
    if haze:
        V, N, F = displaced_sphere(2, seed=44, radius=9.0, center=(0.0, 1.0, 0.0), amplitude=0.0)
        c = sd.add_mesh(V, N, F, bsdf=dict(type="passthrough"))
        sd.set_media(c, interior=hz)
        sd.set_camera_medium(hz)
    V, N, F = displaced_sphere(2, seed=43, radius=0.4, center=(1.5, 0.45, 0.5), amplitude=0.03)
    sd.add_mesh(V, N, F, bsdf=ball_bsdf if ball_bsdf is not None else dict(type="diffuse", reflectance=(0.6, 0.3, 0.2)))
    g = 4.0
    gV = np.array([[-g, 0, -g], [g, 0, -g], [g, 0, g], [-g, 0, g]], np.float32)
    gN = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    sd.add_mesh(gV, gN, np.array([[0, 2, 1], [0, 3, 2]], np.uint32), bsdf=dict(type="diffuse", reflectance=(0.5, 0.5, 0.5)))
    lV, lN, lF = displaced_sphere(1, seed=99, radius=0.4, center=(0.5, 2.6, 0.3), amplitude=0.0)
    sd.add_mesh(lV, lN, lF, bsdf=None, emitter=(25, 22, 18))
    if sphere_light:
        # an area light on an analytic sphere: the scene is no longer "plain", the volumetric kernels run with their emitter /
        # texture / environment-map branches compiled in (render_vol.cuh: PLAIN = false)
        sd.add_sphere((-1.8, 1.6, 1.0), 0.25, bsdf=None, emitter=(12, 14, 18))
    sd.set_sensor(look_at((0.3, 3.0, -4.5), (0, 0.4, 0), (0, 1, 0)), 40.0, width, height, rfilter="box", spp=spp, sampler=sampler, seed=seed,
                  aperture_radius=aperture_radius, focus_distance=5.2)
    sd.set_integrator(kind, max_depth=max_depth, rr_depth=3)
    return sd


def materials_scene(width=48, height=32, spp=4, max_depth=6, sampler="independent", seed=0, subdiv=2, env=True, point=True, materials=None):
    """Displaced spheres, one per entry of MATERIALS, on a diffuse floor; area light (mesh), point light and a
    constant environment — exercises every BSDF and emitter of the GPU subset."""
    sd = SceneDesc()
    if point:
        sd.add_point_light((-3.0, 4.0, -2.0), (20, 18, 15))
    if env:
        sd.add_constant_env((0.3, 0.35, 0.5))
    for i, m in enumerate(MATERIALS if materials is None else materials):
        V, N, F = displaced_sphere(subdiv, seed=20 + i, radius=0.45, center=((i % 4) - 1.5, 0.5, (i // 4) * 1.2 - 0.6))
        c = np.array(((i % 4) - 1.5, 0.5, (i // 4) * 1.2 - 0.6))
        dn = (V - c) / np.linalg.norm(V - c, axis=1, keepdims=True)          # spherical (u,v); required by anisotropic BSDFs
        UV = np.stack([np.arctan2(dn[:, 2], dn[:, 0]) / (2 * math.pi) + 0.5, np.arccos(np.clip(dn[:, 1], -1, 1)) / math.pi], axis=1)
        sd.add_mesh(V, N, F, bsdf=m, UV=UV if i % 2 == 0 else None)
    g = 4.0
    gV = np.array([[-g, 0, -g], [g, 0, -g], [g, 0, g], [-g, 0, g]], np.float32)
    gN = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    sd.add_mesh(gV, gN, np.array([[0, 2, 1], [0, 3, 2]], np.uint32), bsdf=dict(type="diffuse", reflectance=(0.5, 0.5, 0.5)))
    lV, lN, lF = displaced_sphere(1, seed=99, radius=0.4, center=(0.5, 2.6, 0.3), amplitude=0.0)
    sd.add_mesh(lV, lN, lF, bsdf=None, emitter=(25, 22, 18))
    sd.set_sensor(look_at((0.3, 3.0, -4.5), (0, 0.4, 0), (0, 1, 0)), 40.0, width, height, rfilter="box", spp=spp, sampler=sampler, seed=seed)
    sd.set_integrator("path", max_depth=max_depth, rr_depth=3)
    return sd
