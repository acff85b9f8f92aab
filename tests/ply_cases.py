"""Binary PLY file images for the device decoder's parity tests (tests/test_ply_gpu.py) and their golden generator
(tests/golden/make_golden_ply.py): each case is (file image, `ply` shape properties as the reference's XML takes them)."""
import numpy as np


def _grid(n, seed):
    """(n+1)^2 vertices of a bumpy height field, its quads, smooth normals and texture coordinates."""
    rng = np.random.default_rng(seed)
    u, v = np.meshgrid(np.linspace(0, 1, n + 1), np.linspace(0, 1, n + 1), indexing="xy")
    z = 0.15 * np.sin(5 * u) * np.cos(4 * v) + 0.02 * rng.standard_normal(u.shape)
    V = np.stack([2 * u - 1, 2 * v - 1, z], axis=-1).reshape(-1, 3).astype(np.float32)
    N = np.stack([-np.gradient(z, axis=1), -np.gradient(z, axis=0), np.full_like(z, 0.05)], axis=-1).reshape(-1, 3)
    N = (N / np.linalg.norm(N, axis=1, keepdims=True)).astype(np.float32)
    UV = np.stack([u, v], axis=-1).reshape(-1, 2).astype(np.float32)
    idx = np.arange((n + 1) * (n + 1)).reshape(n + 1, n + 1)
    Q = np.stack([idx[:-1, :-1], idx[:-1, 1:], idx[1:, 1:], idx[1:, :-1]], axis=-1).reshape(-1, 4)
    return V, N, UV, Q


def _image(header_lines, vertex_fields, vertex_data, face_dtype, face_data, tail=b"", big_endian=False):
    hdr = "ply\nformat %s 1.0\n" % ("binary_big_endian" if big_endian else "binary_little_endian") + "".join(l + "\n" for l in header_lines) + "end_header\n"
    if big_endian:
        vertex_fields = [(n, t.replace("<", ">")) for n, t in vertex_fields]
        face_dtype = [tuple(x.replace("<", ">") if isinstance(x, str) and i == 1 else x for i, x in enumerate(f)) for f in face_dtype]
    vd = np.empty(len(vertex_data[0]), dtype=vertex_fields)
    for (name, _), col in zip(vertex_fields, vertex_data):
        vd[name] = col
    fd = np.empty(len(face_data[0]), dtype=face_dtype)
    for (name, *_), col in zip(face_dtype, face_data):
        fd[name] = col
    return hdr.encode("ascii") + vd.tobytes() + fd.tobytes() + tail


def tris_of(Q):
    return np.concatenate([Q[:, [0, 1, 2]], Q[:, [0, 2, 3]]], axis=0)


TO_WORLD = '<transform name="toWorld"><scale x="1.5" y="0.7" z="2"/><rotate x="0.3" y="1" z="0.2" angle="37"/><translate x="0.4" y="-1.3" z="2.2"/></transform>'


def to_world_matrix():
    """The same transform as TO_WORLD, composed the way the scene loader does (later entries multiply from the left)."""
    S = np.diag([1.5, 0.7, 2.0, 1.0])
    ax = np.array([0.3, 1.0, 0.2]); ax /= np.linalg.norm(ax)
    a = np.deg2rad(37.0); c, s = np.cos(a), np.sin(a)
    x, y, z = ax
    R = np.eye(4)
    R[:3, :3] = [[c + x*x*(1-c), x*y*(1-c) - z*s, x*z*(1-c) + y*s],
                 [y*x*(1-c) + z*s, c + y*y*(1-c), y*z*(1-c) - x*s],
                 [z*x*(1-c) - y*s, z*y*(1-c) + x*s, c + z*z*(1-c)]]
    T = np.eye(4); T[:3, 3] = [0.4, -1.3, 2.2]
    return T @ R @ S


def cases():
    """name -> dict(image=bytes, xml=extra XML inside <shape type="ply">, kwargs=arguments of Scene.add_ply)"""
    out = {}
    V, N, UV, Q = _grid(12, 1)
    F = tris_of(Q)
    f32 = "<f4"
    # 1. triangles, normals and texture coordinates, identity transform: every array bit-exact
    out["tri_n_uv"] = dict(image=_image(
        ["comment made by tests/ply_cases.py", "element vertex %d" % len(V), "property float x", "property float y", "property float z",
         "property float nx", "property float ny", "property float nz", "property float u", "property float v",
         "element face %d" % len(F), "property list uchar int vertex_indices"],
        [(k, f32) for k in ("x", "y", "z", "nx", "ny", "nz", "u", "v")], [V[:, 0], V[:, 1], V[:, 2], N[:, 0], N[:, 1], N[:, 2], UV[:, 0], UV[:, 1]],
        [("n", "u1"), ("i", "<i4", 3)], [np.full(len(F), 3), F]), xml="", kwargs={})
    # 1b. the same file in binary_big_endian, with a uint count
    out["tri_n_uv_big_endian"] = dict(image=_image(
        ["element vertex %d" % len(V), "property float x", "property float y", "property float z",
         "property float nx", "property float ny", "property float nz", "property float u", "property float v",
         "element face %d" % len(F), "property list uint int vertex_indices"],
        [(k, f32) for k in ("x", "y", "z", "nx", "ny", "nz", "u", "v")], [V[:, 0], V[:, 1], V[:, 2], N[:, 0], N[:, 1], N[:, 2], UV[:, 0], UV[:, 1]],
        [("n", "<u4"), ("i", "<i4", 3)], [np.full(len(F), 3), F], big_endian=True), xml="", kwargs={})
    # 2. no normals in the file: TriMesh::computeNormals makes angle-weighted ones; through toWorld
    out["tri_computed_normals_xform"] = dict(image=_image(
        ["element vertex %d" % len(V), "property float32 x", "property float32 y", "property float32 z",
         "element face %d" % len(F), "property list uint8 int32 vertex_index"],
        [(k, f32) for k in ("x", "y", "z")], [V[:, 0], V[:, 1], V[:, 2]],
        [("n", "u1"), ("i", "<i4", 3)], [np.full(len(F), 3), F]), xml=TO_WORLD, kwargs=dict(to_world=to_world_matrix()))
    # 3. all quads, uint count + uint indices, vertex colours between the fields, a face flag before the list, an element behind
    rgb = (np.random.default_rng(2).random((len(V), 3)) * 255).astype(np.uint8)
    out["quad_colours_extras"] = dict(image=_image(
        ["obj_info something", "element vertex %d" % len(V), "property float x", "property uchar red", "property float y", "property uchar green",
         "property uchar blue", "property float z", "property float s", "property float t",
         "element face %d" % len(Q), "property uchar flag", "property list uint uint vertex_indices",
         "element edge 5", "property int a", "property int b"],
        [("x", f32), ("red", "u1"), ("y", f32), ("green", "u1"), ("blue", "u1"), ("z", f32), ("s", f32), ("t", f32)],
        [V[:, 0], rgb[:, 0], V[:, 1], rgb[:, 1], rgb[:, 2], V[:, 2], UV[:, 0], UV[:, 1]],
        [("flag", "u1"), ("n", "<u4"), ("i", "<u4", 4)], [np.arange(len(Q)) % 7, np.full(len(Q), 4), Q],
        tail=np.arange(10, dtype="<i4").tobytes()), xml="", kwargs={})
    # 4. faceNormals + flipNormals: normals dropped, winding swapped
    out["tri_facenormals_flip"] = dict(image=out["tri_n_uv"]["image"], xml='<boolean name="faceNormals" value="true"/><boolean name="flipNormals" value="true"/>',
                                       kwargs=dict(face_normals=True, flip_normals=True))
    # 5. flipNormals on stored normals, through toWorld (normals go through the inverse transpose)
    out["tri_flip_xform"] = dict(image=out["tri_n_uv"]["image"], xml=TO_WORLD + '<boolean name="flipNormals" value="true"/>',
                                 kwargs=dict(flip_normals=True, to_world=to_world_matrix()))
    # 6. flipNormals on computed normals; a degenerate triangle and an unreferenced vertex (the "bogus" (1,0,0) normal)
    V6 = np.concatenate([V, [[5, 5, 5]]]).astype(np.float32)
    F6 = np.concatenate([F, [[0, 0, 1]]])
    out["tri_computed_flip_degenerate"] = dict(image=_image(
        ["element vertex %d" % len(V6), "property float x", "property float y", "property float z",
         "element face %d" % len(F6), "property list uchar int vertex_indices"],
        [(k, f32) for k in ("x", "y", "z")], [V6[:, 0], V6[:, 1], V6[:, 2]],
        [("n", "u1"), ("i", "<i4", 3)], [np.full(len(F6), 3), F6]), xml='<boolean name="flipNormals" value="true"/>', kwargs=dict(flip_normals=True))
    # 7. triangles and quads mixed in one face element: variable-length records, indexed by one host pass over the count bytes
    nq = len(Q) // 2
    recs = []
    for k, q in enumerate(Q):
        if k < nq or k % 3 == 0:
            recs.append(bytes([4]) + np.asarray(q, "<i4").tobytes())
        else:
            recs.append(bytes([3]) + np.asarray(q[[0, 1, 2]], "<i4").tobytes() + bytes([3]) + np.asarray(q[[0, 2, 3]], "<i4").tobytes())
    nfaces = sum(1 if r[0] == 4 else 2 for r in recs)
    hdr = ("ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
           "element face %d\nproperty list uchar int vertex_indices\nend_header\n" % (len(V), nfaces)).encode("ascii")
    out["mixed_tri_quad_computed"] = dict(image=hdr + np.ascontiguousarray(V, "<f4").tobytes() + b"".join(recs), xml="", kwargs={})
    return out


def ser_cases(scenes):
    """`serialized` file images: name -> dict(image, xml, kwargs of Scene.add_serialized)"""
    out = {}
    V, N, UV, Q = _grid(12, 7)
    F = tris_of(Q)
    V2, N2, UV2, Q2 = _grid(9, 8)
    F2 = tris_of(Q2)
    rgb = np.random.default_rng(4).random((len(V), 3)).astype(np.float32)
    meshes = [dict(V=V, F=F, N=N, UV=UV, name="first"), dict(V=V2 * 0.5, F=F2, name="second, no normals"), dict(V=V, F=F, N=N, C=rgb, name="third")]
    # 1. version 4, single precision, stored normals and uv, no transform: every array bit-exact (normals are NOT re-normalised)
    out["ser_v4_n_uv"] = dict(image=scenes.serialized_image(meshes, 4), xml="", kwargs={})
    # 2. version 4, double precision, second mesh of the file, no normals (computed), through toWorld
    out["ser_v4_double_computed_xform"] = dict(image=scenes.serialized_image(meshes, 4, double=True), xml='<integer name="shapeIndex" value="1"/>' + TO_WORLD,
                                               kwargs=dict(shape_index=1, to_world=to_world_matrix()))
    # 3. version 3, third mesh (colours skipped), mirrored: the winding is swapped, normals go through the inverse transpose
    mirror = np.diag([-1.0, 1.0, 1.0, 1.0])
    out["ser_v3_mirror_xform"] = dict(image=scenes.serialized_image(meshes, 3), xml='<integer name="shapeIndex" value="2"/><transform name="toWorld"><scale x="-1"/></transform>',
                                      kwargs=dict(shape_index=2, to_world=mirror))
    # 4. faceNormals + flipNormals on the first mesh
    out["ser_v4_facenormals_flip"] = dict(image=out["ser_v4_n_uv"]["image"], xml='<boolean name="faceNormals" value="true"/><boolean name="flipNormals" value="true"/>',
                                          kwargs=dict(face_normals=True, flip_normals=True))
    # 5. flipNormals on stored normals, double precision file
    out["ser_v4_double_flip"] = dict(image=out["ser_v4_double_computed_xform"]["image"], xml='<boolean name="flipNormals" value="true"/>', kwargs=dict(flip_normals=True))
    return out


def scene_xml(filename, extra, kind="ply"):
    return ('<?xml version="1.0" encoding="utf-8"?>\n<scene version="0.5.0">\n<shape type="%s"><string name="filename" value="%s"/>%s'
            '<bsdf type="diffuse"/></shape>\n<emitter type="point"><point name="position" x="0" y="0" z="5"/></emitter>\n</scene>\n' % (kind, filename, extra))


def read_mesh_dump(path):
    """oracle/raydump.cpp `mesh` output -> list of (positions, normals or None, texcoords or None, triangles)"""
    raw = np.fromfile(path, np.uint32)
    out, p = [], 0
    while p < raw.size:
        nv, nt, hn, hu = (int(x) for x in raw[p:p + 4]); p += 4
        P = raw[p:p + 3 * nv].view(np.float32).reshape(nv, 3); p += 3 * nv
        N = UV = None
        if hn:
            N = raw[p:p + 3 * nv].view(np.float32).reshape(nv, 3); p += 3 * nv
        if hu:
            UV = raw[p:p + 2 * nv].view(np.float32).reshape(nv, 2); p += 2 * nv
        T = raw[p:p + 3 * nt].reshape(nt, 3); p += 3 * nt
        out.append((P, N, UV, T))
    return out
