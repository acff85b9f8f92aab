"""oracle/oracle.py — TEST INFRASTRUCTURE: ctypes view of oracle/liboracle.so (restate.c) and helpers that run the
UNMODIFIED reference binaries in oracle/_ref.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product path never does."""
import ctypes as C
import json
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            subprocess.check_call(["make", "-C", HERE, "-f", "Makefile"])
        L = C.CDLL(path)
        vp = C.c_void_p
        L.oracle_scene_create.restype = vp
        L.oracle_scene_create.argtypes = [C.c_uint32] + [vp] * 8
        L.oracle_scene_destroy.argtypes = [vp]
        L.oracle_scene_aabb.argtypes = [vp, vp, vp]
        L.oracle_closest.argtypes = [vp, vp, C.c_size_t, vp]
        L.oracle_any.argtypes = [vp, vp, C.c_size_t, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleScene:
    """Brute-force CPU scene built from a scenes.SceneDesc (primitive order = ShapeKDTree::build order)."""

    def __init__(self, desc):
        kinds, vtx_base, tri_base, tri_count, P, T, RW, RO = [], [], [], [], [], [], [], []
        nv = nt = 0
        for s in desc.shapes:
            if s["type"] == "mesh":
                kinds.append(0); vtx_base.append(nv); tri_base.append(nt); tri_count.append(s["F"].shape[0])
                P.append(s["V"].astype(np.float32)); T.append(s["F"].astype(np.uint32))
                nv += s["V"].shape[0]; nt += s["F"].shape[0]
            else:
                kinds.append(1); vtx_base.append(0); tri_base.append(0); tri_count.append(0)
                M = np.asarray(s["to_world"], dtype=np.float64).copy()
                if s.get("flip_normals"):
                    M = M @ np.diag([1.0, 1.0, -1.0, 1.0])
                RW.append(M.astype(np.float32).reshape(-1))
                RO.append(np.linalg.inv(M)[:3, :].astype(np.float32).reshape(-1))
        self._keep = [np.asarray(kinds, np.int32),
                      np.concatenate(P) if P else np.zeros((1, 3), np.float32),
                      np.concatenate(T) if T else np.zeros((1, 3), np.uint32),
                      np.asarray(vtx_base, np.uint32), np.asarray(tri_base, np.uint32), np.asarray(tri_count, np.uint32),
                      np.concatenate(RW) if RW else np.zeros(16, np.float32),
                      np.concatenate(RO) if RO else np.zeros(12, np.float32)]
        k = [np.ascontiguousarray(a) for a in self._keep]
        self._keep = k
        self.h = lib().oracle_scene_create(len(kinds), _p(k[0]), _p(k[1]), _p(k[2]), _p(k[3]), _p(k[4]), _p(k[5]), _p(k[6]), _p(k[7]))

    def aabb(self):
        mn, mx = np.zeros(3, np.float32), np.zeros(3, np.float32)
        lib().oracle_scene_aabb(self.h, _p(mn), _p(mx))
        return mn, mx

    def closest(self, rays):
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        out = np.zeros((r.shape[0], 4), np.float32)
        lib().oracle_closest(self.h, _p(r), r.shape[0], _p(out))
        return out[:, 0].copy(), out[:, 1].copy(), out[:, 2].copy(), out[:, 3].copy().view(np.uint32)

    def any(self, rays):
        r = np.ascontiguousarray(rays).view(np.float32).reshape(-1, 8)
        out = np.zeros(r.shape[0], np.uint8)
        lib().oracle_any(self.h, _p(r), r.shape[0], _p(out))
        return out

    def __del__(self):
        try:
            lib().oracle_scene_destroy(self.h)
        except Exception:
            pass


# --------------------------------------------------------------------------------------------- the real reference
def have_reference():
    return os.path.exists(os.path.join(REF_DIR, "mtsutil")) and os.path.exists(os.path.join(REF_DIR, "plugins", "raydump.so"))


def ref_trace(xml, rays, threads=1, repeats=1, timeout=3600):
    """mtsutil raydump trace (oracle/raydump.cpp) -> (N x 16 float records, timing dict)."""
    with tempfile.TemporaryDirectory() as td:
        rp, op = os.path.join(td, "rays.bin"), os.path.join(td, "out.bin")
        np.ascontiguousarray(rays).view(np.float32).tofile(rp)
        r = subprocess.run([os.path.join(REF_DIR, "mtsutil"), "raydump", "trace", xml, rp, op, str(threads), str(repeats)],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
        if r.returncode != 0 or not os.path.exists(op):
            raise RuntimeError("reference raydump failed:\n" + r.stdout[-3000:])
        out = np.fromfile(op, np.float32).reshape(-1, 16)
    timing = {}
    for line in r.stdout.splitlines():
        if line.startswith("RAYDUMP "):
            timing = json.loads(line[len("RAYDUMP "):])
    return out, timing


def ref_render(xml, out_path, threads=0, classic=True, defines=None, timeout=7200):
    """Runs the reference `mitsuba` CLI; returns (wall seconds of the whole process, render seconds parsed from the log)."""
    import re
    import time
    cmd = [os.path.join(REF_DIR, "mitsuba")]
    if classic:
        cmd.append("-C")
    if threads:
        cmd += ["-p", str(threads)]
    for k, v in (defines or {}).items():
        cmd += ["-D", "%s=%s" % (k, v)]
    cmd += ["-o", out_path, xml]
    t0 = time.time()
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    wall = time.time() - t0
    if r.returncode != 0:
        raise RuntimeError("reference mitsuba failed:\n" + r.stdout[-3000:])
    m = re.search(r"Render time: ([0-9.]+)(ms|s|m)", r.stdout)
    render_s = None
    if m:
        render_s = float(m.group(1)) * {"ms": 1e-3, "s": 1.0, "m": 60.0}[m.group(2)]
    return wall, render_s, r.stdout


def read_pfm(path):
    with open(path, "rb") as f:
        kind = f.readline().strip()
        w, h = (int(x) for x in f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4")
    c = 3 if kind == b"PF" else 1
    return data.reshape(h, w, c)[::-1].astype(np.float32)
