"""Golden vectors for mtsb_eval_transmittance: Scene::evalTransmittance of the UNMODIFIED reference (oracle/raydump.cpp
`transmittance`) on random segments of the fog scene (haze variant: three media, index-matched boundaries, a glass ball, opaque
shapes) -> tests/golden/transmittance_fog.npz.

Run here:  python tests/golden/make_golden_transmittance.py
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

S = g.load_package().scenes
FOG, MILK, HAZE = 0, 1, 2


def medium_at(p):
    """fog ball (index-matched), glass ball of milk, everything else inside the haze ball; -2 = too close to a boundary"""
    out = np.full(len(p), HAZE, np.int32)
    for c, r, m in (((-1.0, 0.65, 0.1), 0.6, FOG), ((0.4, 0.55, -0.2), 0.5, MILK)):
        d = np.linalg.norm(p - np.asarray(c), axis=1)
        out[d < r - 0.08] = m
        out[np.abs(d - r) <= 0.08] = -2
    d = np.linalg.norm(p - np.array([1.5, 0.45, 0.5]), axis=1)      # the diffuse ball: not a place to start from
    out[d < 0.5] = -2
    return out


def segments(n, seed):
    rng = np.random.default_rng(seed)
    lo, hi = np.array([-2.2, 0.05, -1.6]), np.array([2.2, 2.2, 1.6])
    p1 = rng.uniform(lo, hi, (4 * n, 3)); p2 = rng.uniform(lo, hi, (4 * n, 3))
    # a third of the segments start inside the fog ball, a sixth inside the glass ball
    k = 4 * n // 3
    p1[:k] = np.array([-1.0, 0.65, 0.1]) + rng.normal(size=(k, 3)) * 0.25
    p1[k:k + k // 2] = np.array([0.4, 0.55, -0.2]) + rng.normal(size=(k // 2, 3)) * 0.2
    p2[k:k + k // 4] = np.array([0.4, 0.55, -0.2]) + rng.normal(size=(k // 4, 3)) * 0.2
    m = medium_at(p1)
    keep = (m != -2) & (medium_at(p2) != -2)
    p1, p2, m = p1[keep][:n], p2[keep][:n], m[keep][:n]
    flags = rng.integers(0, 4, len(p1))
    # every 16th segment names the wrong medium ("medium inconsistency" when it crosses a boundary), every 16th none at all
    m[::16] = (m[::16] + 1) % 3
    m[8::16] = -1
    return p1.astype(np.float32), p2.astype(np.float32), flags.astype(np.int32), m.astype(np.int32)


def main():
    sd = S.fog_scene(48, 32, spp=4, haze=True)
    p1, p2, flags, med = segments(6000, 17)
    inp = np.zeros((len(p1), 8), np.float32)
    inp[:, 0:3] = p1; inp[:, 3] = flags; inp[:, 4:7] = p2; inp[:, 7] = med
    out = {}
    with tempfile.TemporaryDirectory() as td:
        xml = sd.to_xml(td, "fog")
        ip = os.path.join(td, "in.bin")
        inp.tofile(ip)
        for name, maxi in (("T", -1), ("T_max1", 1)):
            op = os.path.join(td, name + ".bin")
            r = subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", "transmittance", xml, ip, op, str(maxi)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0 or not os.path.exists(op):
                raise RuntimeError(r.stdout[-3000:])
            out[name] = np.fromfile(op, np.float32).reshape(-1, 3)
            print([l for l in r.stdout.splitlines() if l.startswith("transmittance:")])
    T = out["T"]
    print("segments", len(p1), "zero", float((T.max(axis=1) == 0).mean()), "one", float((T.min(axis=1) == 1).mean()),
          "partial", float(((T.max(axis=1) > 0) & (T.min(axis=1) < 1)).mean()), "max1 differs", float((out["T_max1"] != T).any(axis=1).mean()))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "transmittance_fog.npz"), p1=p1, p2=p2, flags=flags, medium=med, **out)


if __name__ == "__main__":
    main()
