"""Golden meshes for the device mesh-file decoders: every file image of tests/ply_cases.py loaded by the UNMODIFIED reference
(`ply` / `serialized` plugins + TriMesh::configure through oracle/raydump.cpp `mesh`) -> tests/golden/ply_meshes.npz.

Run here:  python tests/golden/make_golden_ply.py
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle as O  # noqa: E402
import ply_cases  # noqa: E402


def reference_mesh(image, extra, td, name, kind="ply"):
    with open(os.path.join(td, name + "." + kind), "wb") as f:
        f.write(image)
    xml = os.path.join(td, name + ".xml")
    with open(xml, "w") as f:
        f.write(ply_cases.scene_xml(name + "." + kind, extra, kind))
    out = os.path.join(td, name + ".bin")
    r = subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", "mesh", xml, out], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0 or not os.path.exists(out):
        raise RuntimeError(r.stdout[-3000:])
    meshes = ply_cases.read_mesh_dump(out)
    assert len(meshes) == 1
    return meshes[0]


def main():
    data = {}
    with tempfile.TemporaryDirectory() as td:
        import __graft_entry__ as g
        scenes = g.load_package().scenes
        todo = [(n, c, "ply") for n, c in ply_cases.cases().items()] + [(n, c, "serialized") for n, c in ply_cases.ser_cases(scenes).items()]
        for name, c, kind in todo:
            P, N, UV, T = reference_mesh(c["image"], c["xml"], td, name, kind)
            data[name + "_P"], data[name + "_T"] = P, T
            if N is not None:
                data[name + "_N"] = N
            if UV is not None:
                data[name + "_UV"] = UV
            print(name, P.shape, T.shape, "normals" if N is not None else "-", "uv" if UV is not None else "-")
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ply_meshes.npz"), **data)


if __name__ == "__main__":
    main()
