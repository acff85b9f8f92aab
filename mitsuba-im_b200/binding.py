"""ctypes binding of libmtsb200.so (include/mtsb200.h) — the Python host side used by tests and bench.py.

This is plumbing only: every call forwards to the C ABI; there is no Python or CPU implementation of any
query behind it.  If the shared library is missing the import fails loudly (RuntimeError), and without a
CUDA device every entry point raises MtsbError(MTSB_ENODEVICE).

Names mirror the reference's operators: Scene.ray_intersect  ~ ShapeKDTree::rayIntersect(ray, its)
(src/librender/skdtree.cpp:112-142), Scene.ray_intersect_any ~ ShapeKDTree::rayIntersect(ray) (:207-226),
Render.render_pass ~ SamplingIntegrator::renderBlock (src/librender/integrator.cpp:141-189).
"""
import ctypes as C
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libmtsb200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "mtsb200.h")

MTSB_NO_HIT = 0xFFFFFFFF
PLY_FACE_NORMALS, PLY_FLIP_NORMALS = 1, 2
NOALLOC_MAX_RAYS = 65536
EINVAL, ENODEVICE, ECUDA, EUNSUPPORTED, ENOMEM = -1, -2, -3, -4, -5
SPECIAL_PRIM = 0xF0000000
BSDF_DIFFUSE, BSDF_DIELECTRIC, BSDF_PLASTIC, BSDF_ROUGHCONDUCTOR, BSDF_NULL, BSDF_CONDUCTOR, BSDF_ROUGHDIELECTRIC = 0, 1, 2, 3, 4, 5, 6
BSDF_PASSTHROUGH = 7
BSDF_ROUGHPLASTIC = 8
BSDF_TWOSIDED_FLAG = 0x100
EMITTER_AREA, EMITTER_POINT, EMITTER_CONSTANT, EMITTER_ENVMAP = 0, 1, 2, 3
INTEGRATOR_PATH, INTEGRATOR_DIRECT, INTEGRATOR_VOLPATH_SIMPLE, INTEGRATOR_VOLPATH = 0, 1, 2, 3
BUILDER_LBVH, BUILDER_PLOC = 0, 1

ray_dtype = np.dtype([("o", "<f4", 3), ("mint", "<f4"), ("d", "<f4", 3), ("maxt", "<f4")])
hit_dtype = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim", "<u4")])
surface_dtype = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim", "<u4"), ("shape", "<u4"),
                          ("prim_local", "<u4"), ("p", "<f4", 3), ("ng", "<f4", 3), ("ns", "<f4", 3),
                          ("s", "<f4", 3), ("tt", "<f4", 3), ("uv", "<f4", 2)])
segment_dtype = np.dtype([("p1", "<f4", 3), ("flags", "<u4"), ("p2", "<f4", 3), ("medium", "<i4")])
light_dtype = np.dtype([("t", "<f4"), ("shape", "<u4"), ("n", "<f4", 3), ("uv", "<f4", 2), ("pad", "<u4")])
assert light_dtype.itemsize == 32
assert ray_dtype.itemsize == 32 and hit_dtype.itemsize == 16 and surface_dtype.itemsize == 92


class MtsbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libmtsb200 error %d: %s" % (code, msg))
        self.code = code


class BsdfDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("p", C.c_float * 15)]


class EmitterDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("shape", C.c_uint32), ("radiance", C.c_float * 3), ("pos", C.c_float * 3),
                ("sampling_weight", C.c_float)]


class MediumDesc(C.Structure):
    _fields_ = [("sigma_a", C.c_float * 3), ("sigma_s", C.c_float * 3), ("strategy", C.c_int32),
                ("sampling_density", C.c_float), ("medium_sampling_weight", C.c_float), ("phase", C.c_int32),
                ("g", C.c_float)]


class EnvmapDesc(C.Structure):
    _fields_ = [("levels", C.c_int32), ("sizes", C.POINTER(C.c_int32)), ("data", C.POINTER(C.POINTER(C.c_float))),
                ("scale", C.c_float), ("to_world", C.c_float * 16)]


class TextureDesc(C.Structure):
    _fields_ = [("levels", C.c_int32), ("sizes", C.POINTER(C.c_int32)), ("data", C.POINTER(C.POINTER(C.c_float))),
                ("wrap_u", C.c_int32), ("wrap_v", C.c_int32), ("filter_type", C.c_int32), ("max_anisotropy", C.c_float),
                ("uv_scale", C.c_float * 2), ("uv_offset", C.c_float * 2)]


WRAP_MODES = dict(repeat=0, clamp=1, mirror=2, zero=3, one=4)
FILTER_TYPES = dict(nearest=0, bilinear=1, trilinear=2, ewa=3)


class Camera(C.Structure):
    _fields_ = [("sample_to_camera", C.c_float * 16), ("camera_to_world", C.c_float * 16),
                ("near_clip", C.c_float), ("far_clip", C.c_float), ("inv_resolution", C.c_float * 2),
                ("aperture_radius", C.c_float), ("focus_distance", C.c_float)]


class FilmDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("offset", C.c_int32 * 2), ("border", C.c_int32),
                ("channels", C.c_int32), ("filter_radius", C.c_float), ("filter_table", C.c_float * 32),
                ("crop_offset", C.c_int32 * 2), ("crop_size", C.c_int32 * 2)]


class IntegratorDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("max_depth", C.c_int32), ("rr_depth", C.c_int32), ("strict_normals", C.c_int32),
                ("hide_emitters", C.c_int32), ("emitter_samples", C.c_int32), ("bsdf_samples", C.c_int32)]


def header_symbols():
    """Every function name the C header declares with MTSB_API."""
    text = open(HEADER_PATH).read()
    return sorted(set(re.findall(r"MTSB_API\s+[A-Za-z_ ]+?\*?\s*\*?(mtsb_[a-z0-9_]+)\s*\(", text)))


_lib = None


def lib():
    """Loads libmtsb200.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libmtsb200.so not found at %s — run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, sz, i32, u32, u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint64
    sig = {
        "mtsb_abi_version": (i32, []), "mtsb_last_error": (C.c_char_p, []), "mtsb_device_count": (i32, []),
        "mtsb_ctx_create": (i32, [i32, C.POINTER(vp)]), "mtsb_ctx_destroy": (None, [vp]), "mtsb_ctx_sync": (i32, [vp]),
        "mtsb_ctx_stream": (vp, [vp]),
        "mtsb_scene_create": (i32, [vp, C.POINTER(vp)]), "mtsb_scene_destroy": (None, [vp]),
        "mtsb_scene_add_bsdf": (i32, [vp, C.POINTER(BsdfDesc)]), "mtsb_scene_add_emitter": (i32, [vp, C.POINTER(EmitterDesc)]),
        "mtsb_scene_set_bsdf_table": (i32, [vp, i32, vp, sz]), "mtsb_scene_set_envmap": (i32, [vp, i32, C.POINTER(EnvmapDesc)]),
        "mtsb_scene_add_texture": (i32, [vp, C.POINTER(TextureDesc)]), "mtsb_scene_set_bsdf_texture": (i32, [vp, i32, i32]),
        "mtsb_scene_add_mesh": (i32, [vp, vp, vp, vp, vp, sz, sz, i32, i32]),
        "mtsb_scene_add_ply": (i32, [vp, vp, sz, vp, vp, u32, i32, C.POINTER(u32), C.POINTER(u32)]),
        "mtsb_scene_read_mesh": (i32, [vp, i32, vp, vp, vp, vp]),
        "mtsb_ply_probe": (i32, [vp, sz, C.POINTER(u32), C.POINTER(u32)]),
        "mtsb_scene_add_serialized": (i32, [vp, vp, sz, i32, vp, vp, u32, i32, C.POINTER(u32), C.POINTER(u32)]),
        "mtsb_serialized_probe": (i32, [vp, sz, i32, C.POINTER(u32), C.POINTER(u32)]),
        "mtsb_scene_add_rectangle": (i32, [vp, vp, i32, i32, i32]), "mtsb_scene_add_disk": (i32, [vp, vp, i32, i32, i32]), "mtsb_scene_expand_bounds": (i32, [vp, vp, vp]),
        "mtsb_scene_add_sphere": (i32, [vp, vp, C.c_float, i32, i32, i32]),
        "mtsb_eval_transmittance": (i32, [vp, vp, sz, i32, vp]), "mtsb_scene_add_medium": (i32, [vp, vp]), "mtsb_scene_set_shape_media": (i32, [vp, i32, i32, i32]),
        "mtsb_render_set_camera_medium": (i32, [vp, i32]), "mtsb_render_set_ldsampler": (i32, [vp, i32, sz]), "mtsb_render_sampler_dump": (i32, [vp, u32, u32, u32, u32, u32, vp]), "mtsb_render_set_sobol": (i32, [vp, vp, sz]),
        "mtsb_scene_commit": (i32, [vp, i32]), "mtsb_scene_aabb": (i32, [vp, vp, vp]),
        "mtsb_scene_primitive_count": (sz, [vp]), "mtsb_scene_shape_count": (sz, [vp]),
        "mtsb_scene_shape_offsets": (i32, [vp, vp]), "mtsb_scene_build_stats": (i32, [vp, vp]),
        "mtsb_trace_closest": (i32, [vp, vp, sz, vp, i32]), "mtsb_trace_any": (i32, [vp, vp, sz, vp, i32]),
        "mtsb_trace_surface": (i32, [vp, vp, sz, vp, i32]), "mtsb_trace_light": (i32, [vp, vp, sz, vp, i32]), "mtsb_trace_counters": (i32, [vp, vp]),
        "mtsb_trace_closest_noalloc": (i32, [vp, vp, sz, vp, i32]),
        "mtsb_scene_add_special_disk": (i32, [vp, vp]), "mtsb_trace_closest_all": (i32, [vp, vp, sz, vp, i32]), "mtsb_trace_any_all": (i32, [vp, vp, sz, vp, i32]),
        "mtsb_trace_count": (i32, [vp, vp, sz, i32, i32, vp]),
        "mtsb_host_alloc": (vp, [sz]), "mtsb_host_free": (None, [vp]),
        "mtsb_eval_bsdf": (i32, [vp, i32, vp, sz, vp]), "mtsb_eval_bsdf_s3": (i32, [vp, i32, vp, vp, sz, vp]), "mtsb_eval_emitter": (i32, [vp, vp, sz, vp]),
        "mtsb_render_create": (i32, [vp, C.POINTER(Camera), C.POINTER(FilmDesc), C.POINTER(IntegratorDesc), u64, C.POINTER(vp)]),
        "mtsb_render_destroy": (None, [vp]), "mtsb_render_set_camera": (i32, [vp, C.POINTER(Camera)]),
        "mtsb_render_clear": (i32, [vp]), "mtsb_render_private_stream": (i32, [vp]), "mtsb_film_flush_begin": (i32, [vp]), "mtsb_preview_stack": (i32, [vp, vp, i32, vp, sz]), "mtsb_film_flush_wait": (i32, [vp, vp, vp]), "mtsb_render_reserve": (i32, [vp, u32]), "mtsb_render_set_tail": (i32, [vp, i32]), "mtsb_render_pass": (i32, [vp, u32, u32, u32, vp, u32, u32]),
        "mtsb_render_sync": (i32, [vp]), "mtsb_film_device_ptr": (i32, [vp, C.POINTER(vp), C.POINTER(sz)]),
        "mtsb_film_to_host": (i32, [vp, vp, sz, i32]), "mtsb_film_stage": (i32, [vp, C.POINTER(vp), C.POINTER(sz)]), "mtsb_render_stats": (i32, [vp, vp]),
        "mtsb_render_samples": (i32, [vp, vp, sz, u32, vp]), "mtsb_render_samples_ex": (i32, [vp, vp, sz, u32, vp, vp]),
        "mtsb_comm_version": (i32, []), "mtsb_comm_unique_id": (i32, [vp]),
        "mtsb_comm_create": (i32, [vp, i32, i32, vp, C.POINTER(vp)]), "mtsb_comm_create_all": (i32, [C.POINTER(vp), i32, C.POINTER(vp)]),
        "mtsb_comm_destroy": (None, [vp]), "mtsb_film_reduce": (i32, [vp, vp, i32]),
        "mtsb_film_reduce_all": (i32, [C.POINTER(vp), C.POINTER(vp), i32, i32]),
    }
    for name, (res, args) in sig.items():
        if not hasattr(L, name):
            continue      # symbol table completeness is asserted by tests/test_abi.py
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _check(rc):
    if rc < 0:
        raise MtsbError(rc, lib().mtsb_last_error().decode("utf-8", "replace"))
    return rc


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def ply_probe(image):
    """(vertices, triangles) the device decoder would produce for a binary PLY file image; raises what add_ply would."""
    buf = np.frombuffer(image, dtype=np.uint8)
    nv, nt = C.c_uint32(0), C.c_uint32(0)
    _check(lib().mtsb_ply_probe(_ptr(buf), buf.size, C.byref(nv), C.byref(nt)))
    return nv.value, nt.value


def serialized_probe(image, shape_index=0):
    buf = np.frombuffer(image, dtype=np.uint8)
    nv, nt = C.c_uint32(0), C.c_uint32(0)
    _check(lib().mtsb_serialized_probe(_ptr(buf), buf.size, int(shape_index), C.byref(nv), C.byref(nt)))
    return nv.value, nt.value


def device_count():
    return lib().mtsb_device_count()


class Context:
    """One per GPU (mtsb_ctx)."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        _check(lib().mtsb_ctx_create(device, C.byref(self.h)))
        self.device = device

    def sync(self):
        _check(lib().mtsb_ctx_sync(self.h))

    @property
    def stream(self):
        return lib().mtsb_ctx_stream(self.h)

    def close(self):
        if self.h:
            lib().mtsb_ctx_destroy(self.h)
            self.h = C.c_void_p()


def preview_stack(renders, spp):
    """mtsb_preview_stack: float RGBA image (film size including the border) of the stacked, normalised films."""
    shape = renders[0].film_shape()
    out = np.empty((shape[0], shape[1], 4), np.float32)
    handles = (C.c_void_p * len(renders))(*[r.h for r in renders])
    s = np.ascontiguousarray(spp, dtype=np.float64)
    _check(lib().mtsb_preview_stack(handles, _ptr(s), len(renders), _ptr(out), shape[0] * shape[1]))
    return out


class PinnedArray:
    """numpy view over page-locked memory from mtsb_host_alloc."""

    def __init__(self, shape, dtype):
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        self.ptr = lib().mtsb_host_alloc(max(n, 1))
        if not self.ptr:
            raise MtsbError(-5, "mtsb_host_alloc failed")
        buf = (C.c_uint8 * max(n, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            lib().mtsb_host_free(self.ptr)
            self.ptr = None


class Scene:
    """Device scene (mtsb_scene): shapes are added in Scene::getShapes() order, then commit() builds the BVH."""

    def __init__(self, ctx):
        self.ctx = ctx
        self.h = C.c_void_p()
        _check(lib().mtsb_scene_create(ctx.h, C.byref(self.h)))

    def add_bsdf(self, type_, params=()):
        d = BsdfDesc()
        d.type = type_
        for i, v in enumerate(params):
            d.p[i] = float(v)
        return _check(lib().mtsb_scene_add_bsdf(self.h, C.byref(d)))

    def set_bsdf_table(self, bsdf, table):
        """The rough-transmittance slice of a roughplastic BSDF (hostmath.rough_transmittance_slices)."""
        t = np.ascontiguousarray(table, dtype=np.float32)
        _check(lib().mtsb_scene_set_bsdf_table(self.h, int(bsdf), _ptr(t), t.shape[0]))

    def add_texture(self, pyramid, wrap_u="repeat", wrap_v="repeat", filter_type="ewa", max_anisotropy=20.0, uv_scale=(1, 1), uv_offset=(0, 0)):
        """A `bitmap` texture given by its MIP pyramid (list of [h, w, 3] float arrays, level 0 first); returns its id."""
        levels = [np.ascontiguousarray(a, dtype=np.float32) for a in pyramid]
        d = TextureDesc()
        d.levels = len(levels)
        sizes = (C.c_int32 * (2 * len(levels)))(*[v for a in levels for v in (a.shape[1], a.shape[0])])
        ptrs = (C.POINTER(C.c_float) * len(levels))(*[a.ctypes.data_as(C.POINTER(C.c_float)) for a in levels])
        d.sizes, d.data = sizes, ptrs
        d.wrap_u, d.wrap_v, d.filter_type, d.max_anisotropy = WRAP_MODES[wrap_u], WRAP_MODES[wrap_v], FILTER_TYPES[filter_type], float(max_anisotropy)
        d.uv_scale[:] = [float(uv_scale[0]), float(uv_scale[1])]
        d.uv_offset[:] = [float(uv_offset[0]), float(uv_offset[1])]
        return _check(lib().mtsb_scene_add_texture(self.h, C.byref(d)))

    def set_bsdf_texture(self, bsdf, texture):
        _check(lib().mtsb_scene_set_bsdf_texture(self.h, int(bsdf), int(texture)))

    def set_envmap(self, emitter, pyramid, scale=1.0, to_world=None):
        """The MIP pyramid of an EMITTER_ENVMAP emitter: list of [h, w, 3] float arrays, level 0 first (mtsb_scene_set_envmap)."""
        levels = [np.ascontiguousarray(a, dtype=np.float32) for a in pyramid]
        d = EnvmapDesc()
        d.levels = len(levels)
        sizes = (C.c_int32 * (2 * len(levels)))(*[v for a in levels for v in (a.shape[1], a.shape[0])])
        ptrs = (C.POINTER(C.c_float) * len(levels))(*[a.ctypes.data_as(C.POINTER(C.c_float)) for a in levels])
        d.sizes, d.data, d.scale = sizes, ptrs, float(scale)
        M = np.eye(4, dtype=np.float32) if to_world is None else np.asarray(to_world, np.float32).reshape(4, 4)
        d.to_world[:] = [float(x) for x in M.ravel()]
        _check(lib().mtsb_scene_set_envmap(self.h, int(emitter), C.byref(d)))

    def add_emitter(self, type_, radiance=(1, 1, 1), pos=(0, 0, 0), sampling_weight=1.0):
        d = EmitterDesc()
        d.type = type_
        d.shape = 0
        d.radiance[:] = [float(x) for x in radiance]
        d.pos[:] = [float(x) for x in pos]
        d.sampling_weight = float(sampling_weight)
        return _check(lib().mtsb_scene_add_emitter(self.h, C.byref(d)))

    def add_mesh(self, positions, triangles, normals=None, texcoords=None, bsdf=0, emitter=-1):
        P = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        T = np.ascontiguousarray(triangles, dtype=np.uint32).reshape(-1, 3)
        N = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 3)
        UV = None if texcoords is None else np.ascontiguousarray(texcoords, dtype=np.float32).reshape(-1, 2)
        return _check(lib().mtsb_scene_add_mesh(self.h, _ptr(P), _ptr(N), _ptr(UV), _ptr(T), P.shape[0], T.shape[0], bsdf, emitter))

    def add_ply(self, image, to_world=None, face_normals=False, flip_normals=False, bsdf=0):
        """A binary PLY file image (bytes / numpy uint8 array) decoded on the device at commit; returns the shape index and
        remembers (vertices, triangles) for read_mesh."""
        buf = np.frombuffer(image, dtype=np.uint8) if not isinstance(image, np.ndarray) else np.ascontiguousarray(image, dtype=np.uint8)
        M = Mi = None
        if to_world is not None:
            M = np.ascontiguousarray(to_world, dtype=np.float32).reshape(4, 4)
            Mi = np.ascontiguousarray(np.linalg.inv(M.astype(np.float64)), dtype=np.float32)
        nv, nt = C.c_uint32(0), C.c_uint32(0)
        flags = (PLY_FACE_NORMALS if face_normals else 0) | (PLY_FLIP_NORMALS if flip_normals else 0)
        shape = _check(lib().mtsb_scene_add_ply(self.h, _ptr(buf), buf.size, _ptr(M), _ptr(Mi), flags, bsdf, C.byref(nv), C.byref(nt)))
        if not hasattr(self, "_mesh_sizes"):
            self._mesh_sizes = {}
        self._mesh_sizes[shape] = (nv.value, nt.value)
        return shape

    def add_serialized(self, image, shape_index=0, to_world=None, face_normals=False, flip_normals=False, bsdf=0):
        """A `serialized` mesh file image: inflated on the host, decoded on the device at commit (mtsb_scene_add_serialized)."""
        buf = np.frombuffer(image, dtype=np.uint8) if not isinstance(image, np.ndarray) else np.ascontiguousarray(image, dtype=np.uint8)
        M = Mi = None
        if to_world is not None:
            M = np.ascontiguousarray(to_world, dtype=np.float32).reshape(4, 4)
            Mi = np.ascontiguousarray(np.linalg.inv(M.astype(np.float64)), dtype=np.float32)
        nv, nt = C.c_uint32(0), C.c_uint32(0)
        flags = (PLY_FACE_NORMALS if face_normals else 0) | (PLY_FLIP_NORMALS if flip_normals else 0)
        shape = _check(lib().mtsb_scene_add_serialized(self.h, _ptr(buf), buf.size, int(shape_index), _ptr(M), _ptr(Mi), flags, bsdf, C.byref(nv), C.byref(nt)))
        if not hasattr(self, "_mesh_sizes"):
            self._mesh_sizes = {}
        self._mesh_sizes[shape] = (nv.value, nt.value)
        return shape

    def read_mesh(self, shape, sizes=None):
        """(positions, normals, texcoords, triangles) of a mesh shape as the device holds them (after commit)."""
        nv, nt = sizes if sizes is not None else self._mesh_sizes[shape]
        P, N, UV, T = np.empty((nv, 3), np.float32), np.empty((nv, 3), np.float32), np.empty((nv, 2), np.float32), np.empty((nt, 3), np.uint32)
        _check(lib().mtsb_scene_read_mesh(self.h, shape, _ptr(P), _ptr(N), _ptr(UV), _ptr(T)))
        return P, N, UV, T

    def add_rectangle(self, to_world, flip_normals=False, bsdf=0, emitter=-1):
        M = np.ascontiguousarray(to_world, dtype=np.float32).reshape(4, 4)
        return _check(lib().mtsb_scene_add_rectangle(self.h, _ptr(M), int(flip_normals), bsdf, emitter))

    def expand_bounds(self, lo, hi):
        a, b = np.ascontiguousarray(lo, dtype=np.float32), np.ascontiguousarray(hi, dtype=np.float32)
        _check(lib().mtsb_scene_expand_bounds(self.h, _ptr(a), _ptr(b)))

    def add_disk(self, to_world, flip_normals=False, bsdf=0, emitter=-1):
        M = np.ascontiguousarray(to_world, dtype=np.float32).reshape(4, 4)
        return _check(lib().mtsb_scene_add_disk(self.h, _ptr(M), int(flip_normals), bsdf, emitter))

    def add_sphere(self, object_to_world, radius, flip_normals=False, bsdf=0, emitter=-1):
        M = np.ascontiguousarray(object_to_world, dtype=np.float32).reshape(4, 4)
        return _check(lib().mtsb_scene_add_sphere(self.h, _ptr(M), float(radius), int(flip_normals), bsdf, emitter))

    def add_medium(self, sigma_a, sigma_s, strategy=0, sampling_density=0.0, medium_sampling_weight=None, phase=0, g=0.0):
        """Homogeneous medium; the sampling weight defaults to what the reference's constructor derives
        (src/medium/homogeneous.cpp:171-186): the largest albedo, floored at 0.5 when positive."""
        d = MediumDesc()
        sa, ss = np.asarray(sigma_a, np.float32).reshape(3), np.asarray(sigma_s, np.float32).reshape(3)
        d.sigma_a[:] = [float(x) for x in sa]
        d.sigma_s[:] = [float(x) for x in ss]
        if medium_sampling_weight is None:
            st = sa + ss
            w = 0.0
            for i in range(3):
                if st[i] != 0:
                    w = max(w, float(ss[i] / st[i]))
            medium_sampling_weight = max(w, 0.5) if w > 0 else -1.0
        d.strategy, d.sampling_density, d.medium_sampling_weight = int(strategy), float(sampling_density), float(medium_sampling_weight)
        d.phase, d.g = int(phase), float(g)
        return _check(lib().mtsb_scene_add_medium(self.h, C.byref(d)))

    def set_shape_media(self, shape, interior=-1, exterior=-1):
        _check(lib().mtsb_scene_set_shape_media(self.h, int(shape), int(interior), int(exterior)))

    def commit(self, builder=BUILDER_LBVH):
        _check(lib().mtsb_scene_commit(self.h, builder))

    def aabb(self):
        mn = np.zeros(3, np.float32)
        mx = np.zeros(3, np.float32)
        _check(lib().mtsb_scene_aabb(self.h, _ptr(mn), _ptr(mx)))
        return mn, mx

    @property
    def primitive_count(self):
        return lib().mtsb_scene_primitive_count(self.h)

    def shape_offsets(self):
        n = lib().mtsb_scene_shape_count(self.h)
        out = np.zeros(n + 1, np.uint32)
        _check(lib().mtsb_scene_shape_offsets(self.h, _ptr(out)))
        return out

    def build_stats(self):
        out = np.zeros(8, np.float64)
        _check(lib().mtsb_scene_build_stats(self.h, _ptr(out)))
        return dict(build_ms=out[0], nodes=int(out[1]), node_bytes=int(out[2]), tri_bytes=int(out[3]),
                    sah_cost=out[4], leaves=int(out[5]), depth=int(out[6]), ply_decode_ms=out[7])

    # --- ray queries -------------------------------------------------------------------------
    @staticmethod
    def _rays(rays):
        r = np.ascontiguousarray(rays)
        if r.dtype != ray_dtype:
            r = np.ascontiguousarray(r, dtype=np.float32).reshape(-1, 8).view(ray_dtype).reshape(-1)
        return r

    def ray_intersect(self, rays, out=None):
        """Closest hit of each ray; returns a hit_dtype array (host buffers in, host buffers out)."""
        r = self._rays(rays)
        if out is None:
            out = np.empty(r.shape[0], hit_dtype)
        _check(lib().mtsb_trace_closest(self.h, _ptr(r), r.shape[0], _ptr(out), 0))
        return out

    def ray_intersect_any(self, rays, out=None):
        r = self._rays(rays)
        if out is None:
            out = np.empty(r.shape[0], np.uint8)
        _check(lib().mtsb_trace_any(self.h, _ptr(r), r.shape[0], _ptr(out), 0))
        return out

    def ray_intersect_surface(self, rays):
        r = self._rays(rays)
        out = np.empty(r.shape[0], surface_dtype)
        _check(lib().mtsb_trace_surface(self.h, _ptr(r), r.shape[0], _ptr(out), 0))
        return out

    def eval_transmittance(self, p1, p2, medium, p1_on_surface=False, p2_on_surface=False, max_interactions=-1):
        """Scene::evalTransmittance for a batch of segments (mtsb_eval_transmittance); returns [n, 3]."""
        p1 = np.asarray(p1, np.float32).reshape(-1, 3)
        seg = np.zeros(p1.shape[0], segment_dtype)
        seg["p1"] = p1; seg["p2"] = np.asarray(p2, np.float32).reshape(-1, 3)
        seg["medium"] = medium
        seg["flags"] = np.asarray(p1_on_surface, np.uint32) * 1 + np.asarray(p2_on_surface, np.uint32) * 2
        return self.eval_transmittance_segments(seg, max_interactions=max_interactions)

    def eval_transmittance_segments(self, segments, out=None, max_interactions=-1):
        """mtsb_eval_transmittance on a prepared segment_dtype array (what a C caller hands over)."""
        seg = np.ascontiguousarray(segments, dtype=segment_dtype)
        if out is None:
            out = np.zeros((seg.shape[0], 3), np.float32)
        _check(lib().mtsb_eval_transmittance(self.h, _ptr(seg), seg.shape[0], int(max_interactions), _ptr(out)))
        return out

    def ray_intersect_light(self, rays):
        """ShapeKDTree::rayIntersect(ray, t, shape, n, uv) (src/librender/skdtree.cpp:144-204) for a batch."""
        r = self._rays(rays)
        out = np.empty(r.shape[0], light_dtype)
        _check(lib().mtsb_trace_light(self.h, _ptr(r), r.shape[0], _ptr(out), 0))
        return out

    def trace_device(self, mode, d_rays, n, d_out):
        """Asynchronous query on device pointers (ints); mode in {'closest','any','surface'}."""
        fn = {"closest": lib().mtsb_trace_closest, "any": lib().mtsb_trace_any, "surface": lib().mtsb_trace_surface}[mode]
        _check(fn(self.h, C.c_void_p(d_rays), n, C.c_void_p(d_out), 1))

    def add_special_disk(self, to_world):
        """A special shape of Scene::initializeBidirectional outside the kd-tree: the aperture disk of a thinlens sensor."""
        M = np.ascontiguousarray(to_world, dtype=np.float32).reshape(4, 4)
        return _check(lib().mtsb_scene_add_special_disk(self.h, _ptr(M)))

    def ray_intersect_all(self, rays):
        """Scene::rayIntersectAll(ray, its): closest hit over the kd-tree shapes and the special shapes (prim >= SPECIAL_PRIM)."""
        r = self._rays(rays)
        out = np.empty(r.shape[0], hit_dtype)
        _check(lib().mtsb_trace_closest_all(self.h, _ptr(r), r.shape[0], _ptr(out), 0))
        return out

    def ray_intersect_any_all(self, rays):
        r = self._rays(rays)
        out = np.empty(r.shape[0], np.uint8)
        _check(lib().mtsb_trace_any_all(self.h, _ptr(r), r.shape[0], _ptr(out), 0))
        return out

    def ray_intersect_noalloc(self, rays, out=None):
        """mtsb_trace_closest_noalloc with host buffers: at most NOALLOC_MAX_RAYS rays, no device allocation inside."""
        r = self._rays(rays)
        if out is None:
            out = np.empty(r.shape[0], hit_dtype)
        _check(lib().mtsb_trace_closest_noalloc(self.h, _ptr(r), r.shape[0], _ptr(out), 0))
        return out

    def trace_device_noalloc(self, d_rays, n, d_hits):
        """mtsb_trace_closest_noalloc on device pointers (ints); asynchronous on the context stream."""
        _check(lib().mtsb_trace_closest_noalloc(self.h, C.c_void_p(d_rays), n, C.c_void_p(d_hits), 1))

    def trace_count(self, rays, any_hit=False):
        r = self._rays(rays)
        out = np.zeros(4, np.uint64)
        _check(lib().mtsb_trace_count(self.h, _ptr(r), r.shape[0], int(any_hit), 0, _ptr(out)))
        return dict(nodes=int(out[0]), prims=int(out[1]), rays=int(out[2]))

    def trace_count_device(self, d_rays, n, any_hit=False):
        out = np.zeros(4, np.uint64)
        _check(lib().mtsb_trace_count(self.h, C.c_void_p(d_rays), n, int(any_hit), 1, _ptr(out)))
        return dict(nodes=int(out[0]), prims=int(out[1]), rays=int(out[2]))

    # --- pointwise plugin math ------------------------------------------------------------------
    def eval_bsdf(self, bsdf, wi, wo, samples, s3=None):
        n = len(wi)
        inp = np.zeros((n, 8), np.float32)
        inp[:, 0:3] = wi
        inp[:, 3:6] = wo
        inp[:, 6:8] = samples
        out = np.zeros((n, 16), np.float32)
        if s3 is not None:
            s3 = np.ascontiguousarray(s3, np.float32)
            _check(lib().mtsb_eval_bsdf_s3(self.h, bsdf, _ptr(inp), _ptr(s3), n, _ptr(out)))
        else:
            _check(lib().mtsb_eval_bsdf(self.h, bsdf, _ptr(inp), n, _ptr(out)))
        return out

    def eval_emitter(self, p, n_ref, samples):
        n = len(p)
        inp = np.zeros((n, 8), np.float32)
        inp[:, 0:3] = p
        inp[:, 3:6] = n_ref
        inp[:, 6:8] = samples
        out = np.zeros((n, 12), np.float32)
        _check(lib().mtsb_eval_emitter(self.h, _ptr(inp), n, _ptr(out)))
        return out

    def close(self):
        if self.h:
            lib().mtsb_scene_destroy(self.h)
            self.h = C.c_void_p()


class Render:
    """Wavefront renderer state (mtsb_render): device film in ImageBlock layout + path queues."""

    def __init__(self, scene, camera, film, integrator, seed=0):
        self.scene = scene
        self.film = film
        self.h = C.c_void_p()
        _check(lib().mtsb_render_create(scene.h, C.byref(camera), C.byref(film), C.byref(integrator), seed, C.byref(self.h)))

    def set_camera(self, camera):
        _check(lib().mtsb_render_set_camera(self.h, C.byref(camera)))

    def set_sobol(self, tables):
        """Switches to the `sobol` sampler; tables = the reference's direction numbers as uint32 words (mtsb_render_set_sobol)."""
        t = np.ascontiguousarray(tables, dtype=np.uint32)
        _check(lib().mtsb_render_set_sobol(self.h, _ptr(t), t.shape[0]))

    def set_ldsampler(self, dimension, sample_count):
        """Switches to the `ldsampler` ((0,2)-sequence point sets per pixel, mtsb_render_set_ldsampler)."""
        _check(lib().mtsb_render_set_ldsampler(self.h, int(dimension), int(sample_count)))

    def sampler_dump(self, px, py, first, n, draws):
        """[n, draws, 3] = (next2D.x, next2D.y, next1D) the sampler hands to samples first..first+n-1 of a pixel (test support)."""
        out = np.zeros((n, draws, 3), np.float32)
        _check(lib().mtsb_render_sampler_dump(self.h, px, py, first, n, draws, _ptr(out)))
        return out

    def set_camera_medium(self, medium):
        _check(lib().mtsb_render_set_camera_medium(self.h, int(medium)))

    def clear(self):
        _check(lib().mtsb_render_clear(self.h))

    def private_stream(self):
        _check(lib().mtsb_render_private_stream(self.h))

    def reserve(self, spp_per_pass):
        _check(lib().mtsb_render_reserve(self.h, spp_per_pass))

    def set_tail(self, paths):
        """Live-path count below which the late-bounce kernel takes over (0 = wavefront stages only, -1 = default)."""
        _check(lib().mtsb_render_set_tail(self.h, int(paths)))

    def render_pass(self, spp_begin, spp_count, spp_total, rect=None, stride=1, phase=0):
        r = None if rect is None else np.ascontiguousarray(rect, dtype=np.int32)
        _check(lib().mtsb_render_pass(self.h, spp_begin, spp_count, spp_total, _ptr(r), stride, phase))

    def sync(self):
        _check(lib().mtsb_render_sync(self.h))

    def film_shape(self):
        f = self.film
        return (f.height + 2 * f.border, f.width + 2 * f.border, f.channels)

    def film_device_ptr(self):
        p = C.c_void_p()
        n = C.c_size_t()
        _check(lib().mtsb_film_device_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def film_to_host(self, out=None, accumulate=False):
        if out is None:
            out = np.zeros(self.film_shape(), np.float32)
        _check(lib().mtsb_film_to_host(self.h, _ptr(out), out.size, int(accumulate)))
        return out

    def stats(self):
        out = np.zeros(8, np.uint64)
        _check(lib().mtsb_render_stats(self.h, _ptr(out)))
        return dict(paths=int(out[0]), closest_rays=int(out[1]), shadow_rays=int(out[2]), depth_sum=int(out[3]),
                    launches=int(out[4]))

    def render_samples(self, samples, spp_total, positions=False):
        s = np.ascontiguousarray(samples, dtype=np.uint32).reshape(-1, 4)
        out = np.zeros((s.shape[0], 4), np.float32)
        if not positions:
            _check(lib().mtsb_render_samples(self.h, _ptr(s), s.shape[0], spp_total, _ptr(out)))
            return out
        pos = np.zeros((s.shape[0], 2), np.float32)
        _check(lib().mtsb_render_samples_ex(self.h, _ptr(s), s.shape[0], spp_total, _ptr(out), _ptr(pos)))
        return out, pos

    def close(self):
        if self.h:
            lib().mtsb_render_destroy(self.h)
            self.h = C.c_void_p()


COMM_ID_BYTES = 128


def comm_unique_id():
    """ncclUniqueId bytes made by rank 0 and shipped to the other ranks by the launcher's side channel."""
    buf = (C.c_uint8 * COMM_ID_BYTES)()
    _check(lib().mtsb_comm_unique_id(buf))
    return bytes(buf)


class Comm:
    """NCCL communicator of one rank (mtsb_comm), used only for the film reduction at the end of a render."""

    def __init__(self, ctx, nranks, rank, unique_id):
        self.h = C.c_void_p()
        self.nranks, self.rank = nranks, rank
        buf = (C.c_uint8 * COMM_ID_BYTES).from_buffer_copy(unique_id)
        _check(lib().mtsb_comm_create(ctx.h, nranks, rank, buf, C.byref(self.h)))

    def film_reduce(self, render, root=0):
        """film(root) += films of every other rank; asynchronous on the context stream."""
        _check(lib().mtsb_film_reduce(render.h, self.h, root))

    def close(self):
        if self.h:
            lib().mtsb_comm_destroy(self.h)
            self.h = C.c_void_p()


def comm_create_all(contexts):
    """Single process, one context per GPU: list of Comm objects (rank i on contexts[i])."""
    n = len(contexts)
    ctxs = (C.c_void_p * n)(*[c.h for c in contexts])
    out = (C.c_void_p * n)()
    _check(lib().mtsb_comm_create_all(ctxs, n, out))
    comms = []
    for i in range(n):
        c = Comm.__new__(Comm)
        c.h, c.nranks, c.rank = C.c_void_p(out[i]), n, i
        comms.append(c)
    return comms


def film_reduce_all(renders, comms, root=0):
    n = len(renders)
    r = (C.c_void_p * n)(*[x.h for x in renders])
    c = (C.c_void_p * n)(*[x.h for x in comms])
    _check(lib().mtsb_film_reduce_all(r, c, n, root))
