"""Goldens for the `ldsampler`, from the UNMODIFIED reference:

  qmc.npz                 radicalInverse2Single / sobol2Single (include/mitsuba/core/qmc.h) on random (n, scramble) words — pins the
                          numpy restatement in tests/ld_restate.py, which pins the device's point sets bit for bit
  li_ldsampler_cornell.npz   the Cornell box (64 x 64, path, maxDepth 5) rendered by the reference with `ldsampler` at 16 spp (four
                          runs: the sampler draws from a sequential Random, every run is another realisation), with `independent` at
                          16 spp (four runs) and at 4096 spp (the converged image the errors are measured against)

Run here:  python tests/golden/make_golden_ld.py
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
GOLD = os.path.join(ROOT, "tests", "golden")


def qmc():
    rng = np.random.default_rng(11)
    n = np.concatenate([np.arange(64, dtype=np.uint32), rng.integers(0, 1 << 24, 4000, dtype=np.uint32), rng.integers(0, 1 << 32, 4000, dtype=np.uint64).astype(np.uint32)])
    s0 = rng.integers(0, 1 << 32, n.size, dtype=np.uint64).astype(np.uint32)
    s1 = rng.integers(0, 1 << 32, n.size, dtype=np.uint64).astype(np.uint32)
    s0[:64] = 0; s1[:64] = 0
    with tempfile.TemporaryDirectory() as td:
        ip, op = os.path.join(td, "in.bin"), os.path.join(td, "out.bin")
        np.stack([n, s0, s1], axis=1).astype(np.uint32).tofile(ip)
        r = subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", "qmc", ip, op], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(r.stdout[-2000:])
        out = np.fromfile(op, np.float32).reshape(-1, 2)
    np.savez_compressed(os.path.join(GOLD, "qmc.npz"), n=n, s0=s0, s1=s1, radical_inverse=out[:, 0], sobol2=out[:, 1])
    print("qmc", out.shape)


def images():
    res = {}
    with tempfile.TemporaryDirectory() as td:
        def render(kind, spp, name, threads=4):
            sd = S.cornell_box(64, 64, spp=spp, max_depth=5, sampler=kind, rfilter="box")
            xml = sd.to_xml(td, name)
            O.ref_render(xml, os.path.join(td, name + ".pfm"), threads=threads, classic=True)
            return O.read_pfm(os.path.join(td, name + ".pfm"))[..., :3].astype(np.float32)
        res["ref_hi"] = render("independent", 4096, "hi", threads=0)
        res["ref_ld"] = np.stack([render("ldsampler", 16, "ld%d" % k) for k in range(4)])
        res["ref_ind"] = np.stack([render("independent", 16, "ind%d" % k) for k in range(4)])
    hi = res["ref_hi"]
    rel = lambda img: float(np.mean((img - hi) ** 2 / (hi ** 2 + 1e-3)))
    print("relMSE ldsampler", [round(rel(i), 5) for i in res["ref_ld"]], "independent", [round(rel(i), 5) for i in res["ref_ind"]])
    np.savez_compressed(os.path.join(GOLD, "li_ldsampler_cornell.npz"), **res)


if __name__ == "__main__":
    qmc()
    images()
