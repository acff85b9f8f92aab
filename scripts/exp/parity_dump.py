"""Diagnostic (GPU box): dumps the rays on which the device and the reference disagree, for the config-2 shadow set and the
config-3 sets, with both sides' records -> gpurun_out/parity_dump.npz (analysed offline with the regenerated meshes)."""
import importlib.util, os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as graft
import oracle as O
pkg = graft.load_package(); S, B = pkg.scenes, pkg.binding
ctx = B.Context(0)
out = {}

def dump(tag, sc, rays, ref_t, ref_prim, ref_u, ref_v, ref_any):
    h = sc.ray_intersect_surface(rays)
    a = sc.ray_intersect_any(rays)
    prim = h["prim"].view(np.int32)
    hit, rhit = np.isfinite(h["t"]), np.isfinite(ref_t)
    bad = (hit != rhit) | (rhit & hit & (prim != ref_prim)) | (a.astype(bool) != ref_any.astype(bool))
    idx = np.nonzero(bad)[0]
    print(tag, "n", len(rays), "disagree", len(idx))
    out[tag + "_rays"] = np.asarray(rays).view(np.float32).reshape(-1, 8)[idx]
    out[tag + "_ours"] = np.stack([h["t"][idx], h["u"][idx], h["v"][idx], prim[idx].astype(np.float32)], axis=1)
    out[tag + "_ours_prim"] = prim[idx]
    out[tag + "_ours_any"] = a[idx]
    out[tag + "_ref"] = np.stack([ref_t[idx], ref_u[idx], ref_v[idx]], axis=1)
    out[tag + "_ref_prim"] = ref_prim[idx]
    out[tag + "_ref_any"] = ref_any[idx]
    out[tag + "_idx"] = idx

# config 2
V, N, F = S.displaced_sphere(8, seed=1)
sd = S.SceneDesc(); sd.add_mesh(V, N, F); sd.add_point_light((0, 5, 0), (1, 1, 1))
td = tempfile.mkdtemp()
xml = sd.to_xml(td, "c2")
sc = sd.upload(ctx, B.BUILDER_LBVH)
coh = S.pinhole_rays(1024, origin=(0, 0.3, -3.0))
h1 = sc.ray_intersect_surface(coh)
sh = S.shadow_rays(h1["p"][np.isfinite(h1["t"])], light_center=(0, 3.0, 0.5))
ref, timing = O.ref_trace(xml, sh, threads=os.cpu_count() or 1)
ref_t = ref[:, 0].copy(); ref_prim = np.where(np.isfinite(ref_t), ref[:, 4].copy().view(np.int32), -1)
dump("c2shadow", sc, sh, ref_t, ref_prim, ref[:, 1].copy(), ref[:, 2].copy(), ref[:, 14].copy().view(np.int32))
sc.close()

# config 3 (golden)
spec = importlib.util.spec_from_file_location("mk", os.path.join(ROOT, "tests", "golden", "make_golden_c3.py"))
mk = importlib.util.module_from_spec(spec); spec.loader.exec_module(mk)
gold = np.load(os.path.join(ROOT, "tests", "golden", "hits_c3.npz"))
sd3 = S.sphere_field()
n = int(gold["n"]); sizes = gold["sizes"]
r0, r1 = mk.c3_primary(S, n)
off = np.cumsum(np.concatenate([[0], sizes]))
r2, r3 = mk.c3_derived(S, r1, gold["ref_t"][off[1]:off[2]])
ref_any = np.unpackbits(gold["ref_any"])[: off[-1]]
sc = sd3.upload(ctx, B.BUILDER_PLOC)
for k, (name, rays) in enumerate((("chord", r0), ("primary", r1), ("secondary", r2), ("shadow", r3))):
    a, b = off[k], off[k + 1]
    dump("c3" + name, sc, rays, gold["ref_t"][a:b], gold["ref_prim"][a:b], gold["ref_u"][a:b], gold["ref_v"][a:b], ref_any[a:b])
sc.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "parity_dump.npz"), **out)
