"""Three-way comparison on the haze scene: stock volpath_simple, the overlay plugin, the C ABI (same ctr samples)."""
import os, sys, subprocess
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g
import oracle as O
pkg = g.load_package()
S, B = pkg.scenes, pkg.binding
OVERLAY = os.path.join(ROOT, "mitsuba-im_b200", "plugin", "overlay")
out = os.path.join(ROOT, "gpurun_out"); os.makedirs(out, exist_ok=True)
haze = len(sys.argv) < 2 or sys.argv[1] != "nohaze"
sd = S.fog_scene(64, 48, spp=8, max_depth=8, sampler="ctr", seed=53, haze=haze)
xml = sd.to_xml(out, "fogdbg")
def run(name, overlay, classic=True, xml=xml):
    cmd = [os.path.join(O.REF_DIR, "mitsuba"), "-p", "4"] + (["-C"] if classic else []) + (["-a", OVERLAY] if overlay else []) + ["-o", os.path.join(out, name), xml]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, errors="replace")
    open(os.path.join(out, name + ".log"), "w").write(r.stdout)
    return O.read_pfm(os.path.join(out, name))
ref = run("dbg_ref.pfm", False)
gpu = run("dbg_gpu.pfm", True)
gpu_r = run("dbg_gpu_r.pfm", True, classic=False)
ref_r = run("dbg_ref_r.pfm", False, classic=False)
print("responsive: gpu mean", gpu_r.mean(), "classic gpu mean", gpu.mean(), "ref responsive mean", ref_r.mean(), "ref classic", ref.mean())
rr = np.abs(gpu_r - gpu).mean(axis=(1, 2)) / gpu.mean()
print("per-row |responsive - classic| / mean:", np.round(rr, 3))
sdp = S.fog_scene(64, 48, spp=8, max_depth=8, sampler="ctr", seed=53, haze=haze, kind="path")
xmlp = sdp.to_xml(out, "fogdbg_path")
pc = run("dbg_pc.pfm", True, True, xmlp); pr = run("dbg_pr.pfm", True, False, xmlp)
print("path integrator on the same scene: classic mean", pc.mean(), "responsive mean", pr.mean())
ctx = B.Context(0)
sc = sd.upload(ctx, B.BUILDER_PLOC)
rd = sd.render(sc)
w, h, spp = 64, 48, 8
ys, xs, ss = np.meshgrid(np.arange(h), np.arange(w), np.arange(spp), indexing="ij")
samples = np.stack([xs.ravel(), ys.ravel(), ss.ravel(), np.zeros(xs.size)], axis=1).astype(np.uint32)
L = rd.render_samples(samples, spp).reshape(h, w, spp, 4)
abi = L[..., :3].mean(axis=2)
def frac(a, b): return float((np.abs(a - b) <= 2e-4 + 2e-3 * np.abs(b)).all(axis=2).mean())
print("plugin vs ref", frac(gpu, ref), "abi vs ref", frac(abi, ref), "plugin vs abi", frac(gpu, abi))
rows = (np.abs(gpu - ref) <= 2e-4 + 2e-3 * np.abs(ref)).all(axis=2).mean(axis=1)
print("per-row agreement plugin/ref:", np.round(rows, 2))
print("row 2 ref", ref[2, :4], "gpu", gpu[2, :4], "abi", abi[2, :4])
print("ratio gpu/ref sky mean", (gpu[:6] / ref[:6]).mean(axis=(0, 1)))
print(open(os.path.join(out, "dbg_gpu.pfm.log")).read()[-1500:])
