"""SURVEY.md §8 f1 — the per-ray drop-in: with libmtsb_overlay.so pre-loaded, the three ShapeKDTree::rayIntersect overloads
of the UNMODIFIED reference (src/librender/skdtree.cpp:112-226) are served by the GPU for every caller in the process.  The
reference's own CPU integrators (`path`, `direct` — stock plugins, no GPU integrator involved) then render the same image
as without it: same random numbers (ctr sampler), same hits, so the same paths."""
import os
import re
import subprocess
import time

import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PRELOAD = os.path.join(ROOT, "mitsuba-im_b200", "plugin", "overlay", "lib", "libmtsb_overlay.so")
XML = os.path.join(ROOT, "tests", "golden", "cbox_config1.xml")


def run_stock(out, defines, threads, preload, xml=XML):
    env = dict(os.environ)
    if preload:
        env["LD_PRELOAD"] = PRELOAD
    cmd = [os.path.join(O.REF_DIR, "mitsuba"), "-C", "-p", str(threads)]
    for k, v in defines.items():
        cmd += ["-D", "%s=%s" % (k, v)]
    cmd += ["-o", out, xml]
    t0 = time.perf_counter()
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1800, env=env)
    wall = time.perf_counter() - t0
    assert r.returncode == 0, r.stdout[-3000:]
    m = re.search(r"Render time: ([0-9.]+)(ms|s|m)", r.stdout)
    render_s = float(m.group(1)) * {"ms": 1e-3, "s": 1.0, "m": 60.0}[m.group(2)] if m else wall
    return O.read_pfm(out), r.stdout, render_s


@pytest.mark.parametrize("integrator", ["path", "direct"])
def test_stock_cpu_integrators_through_the_overlay(tmp_path, integrator):
    if not O.have_reference() or not os.path.exists(PRELOAD):
        pytest.fail("oracle/_ref or libmtsb_overlay.so is missing: run __graft_entry__.build() before shipping")
    xml = XML
    if integrator == "direct":
        xml = str(tmp_path / "cbox_direct.xml")
        open(xml, "w").write(open(XML).read().replace('<integrator type="path"><integer name="maxDepth" value="5"/></integrator>',
                                                      '<integrator type="direct"><integer name="shadingSamples" value="2"/></integrator>'))
    d = dict(sampler="ctr", spp=4, res=64, rfilter="box")
    ref, _, t_ref = run_stock(str(tmp_path / "ref.pfm"), d, 8, False, xml)
    gpu, log, t_gpu = run_stock(str(tmp_path / "gpu.pfm"), d, 128, True, xml)
    assert "mtsb overlay: ShapeKDTree::rayIntersect x3 now run on the GPU" in log, log[-2000:]
    assert "GPU integrator" not in log and "gpupath" not in log         # the integrator is the reference's own CPU plugin
    batches = re.findall(r"mtsb overlay: (\w+) queue: (\d+) rays in (\d+) batches", log)
    print("overlay (%s): stock %.3f s, through the GPU %.3f s; queues %s" % (integrator, t_ref, t_gpu, batches))
    assert batches and sum(int(b[1]) for b in batches) > 64 * 64 * 4
    ok = (np.abs(gpu - ref) <= 2e-4 + 2e-3 * np.abs(ref)).all(axis=2)
    assert ok.mean() >= 0.97, "%.4f of the pixels match the stock render" % ok.mean()
    assert abs(gpu.mean() - ref.mean()) <= 0.01 * ref.mean()


def test_overlay_skips_the_cpu_kdtree_build(pkg, tmp_path):
    """SURVEY.md §8 f3 through the overlay: with the queries on the GPU the interposed ShapeKDTree::build() keeps the prefix
    sums and the boxes of skdtree.cpp:68-70 / gkdtree.h:997-1002,1213-1220 and leaves out the SAH build.  A 225 K-triangle
    scene (meshes, point / area / environment emitters — the environment's sphere and the sensor's clip planes hang on the
    boxes) renders the same image through the stock CPU `path` plugin, and through the GPU integrator plugin on top of it;
    MTSB_OVERLAY_KDBUILD=1 brings the reference's build back."""
    OVERLAY = os.path.join(ROOT, "mitsuba-im_b200", "plugin", "overlay")
    sd = pkg.scenes.materials_scene(64, 48, spp=4, max_depth=5, sampler="ctr", seed=61, subdiv=5)
    xml = sd.to_xml(str(tmp_path), "big")
    ref, log_ref, _ = run_stock(str(tmp_path / "ref.pfm"), {}, 8, False, xml)
    assert "Constructing a SAH kd-tree" in log_ref
    gpu, log, _ = run_stock(str(tmp_path / "gpu.pfm"), {}, 128, True, xml)
    assert "the CPU kd-tree is not built" in log and "Constructing a SAH kd-tree" not in log, log[-3000:]
    assert "225280 primitives" in log or "primitives" in log
    ok = (np.abs(gpu - ref) <= 2e-4 + 2e-3 * np.abs(ref)).all(axis=2)
    assert ok.mean() >= 0.97, "%.4f of the pixels match the stock render" % ok.mean()
    # the reference's build on request
    env = dict(os.environ, LD_PRELOAD=PRELOAD, MTSB_OVERLAY_KDBUILD="1")
    r = subprocess.run([os.path.join(O.REF_DIR, "mitsuba"), "-C", "-p", "64", "-o", str(tmp_path / "kd.pfm"), xml],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1800, env=env)
    assert r.returncode == 0 and "Constructing a SAH kd-tree" in r.stdout and "the CPU kd-tree is not built" not in r.stdout, r.stdout[-3000:]
    # the GPU integrator plugin (overlay plugin directory) on top of the pre-loaded library: no kd-tree, same image
    env = dict(os.environ, LD_PRELOAD=PRELOAD)
    t0 = time.perf_counter()
    r = subprocess.run([os.path.join(O.REF_DIR, "mitsuba"), "-C", "-p", "4", "-a", OVERLAY, "-o", str(tmp_path / "both.pfm"), xml],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1800, env=env)
    assert r.returncode == 0 and "gpupath:" in r.stdout and "the CPU kd-tree is not built" in r.stdout, r.stdout[-3000:]
    both = O.read_pfm(str(tmp_path / "both.pfm"))
    ok = (np.abs(both - ref) <= 2e-4 + 2e-3 * np.abs(ref)).all(axis=2)
    assert ok.mean() >= 0.96, "%.4f of the pixels match the stock render" % ok.mean()
    print("overlay + gpupath, no kd-tree: process wall %.2f s" % (time.perf_counter() - t0))
