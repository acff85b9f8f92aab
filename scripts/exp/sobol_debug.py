import os, sys, subprocess
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as g, oracle as O
pkg = g.load_package(); S, B = pkg.scenes, pkg.binding
out = "/tmp/sobol.bin"
subprocess.run([os.path.join(O.REF_DIR, "mtsutil"), "raydump", "sobol", out], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
S.SOBOL_TABLES = np.fromfile(out, np.uint32)
gold = np.load(os.path.join(ROOT, "tests/golden/li_sobol_cornell.npz"))
ref, spp = gold["image"], int(gold["spp"])
sd = S.cornell_box(40, 24, spp=spp, max_depth=5, sampler="sobol", seed=0)
ctx = B.Context(0); sc = sd.upload(ctx); rd = sd.render(sc)
h, w = 24, 40
ys, xs, ss = np.meshgrid(np.arange(h), np.arange(w), np.arange(spp), indexing="ij")
samples = np.stack([xs.ravel(), ys.ravel(), ss.ravel(), np.zeros(xs.size)], axis=1).astype(np.uint32)
L, pos = rd.render_samples(samples, spp, positions=True)
L = L.reshape(h, w, spp, 4); pos = pos.reshape(h, w, spp, 2)
img = L[..., :3].mean(axis=2)
ok = (np.abs(img - ref) <= 2e-4 + 2e-3 * np.abs(ref)).all(axis=2)
print("agree", ok.mean())
bad = np.argwhere(~ok)
print("bad pixels (y,x):", bad[:40].tolist())
for y, x in bad[:6]:
    print((y, x), "img", img[y, x], "ref", ref[y, x], "positions", pos[y, x].tolist())
print("rows with failures:", np.unique(bad[:, 0]), "cols:", np.unique(bad[:, 1]))
per = np.load(os.path.join(ROOT, "scripts/exp/tmp_sobol_ref.npz"))["per"]     # [spp, h, w, 3] per-sample reference images (differences of means)
for k in range(spp):
    okk = (np.abs(L[:, :, k, :3] - per[k]) <= 1e-3 + 1e-2 * np.abs(per[k])).all(axis=2)
    print("sample", k, "agree", okk.mean(), "bad:", np.argwhere(~okk)[:12].tolist())
for (y, x) in ((1, 33), (2, 16), (4, 14)):
    for k in range(4):
        print((y, x), "sample", k, "gpu", L[y, x, k, :3], "ref", per[k, y, x])
# same scene, independent (ctr) sampler through both: which pixels differ there?
