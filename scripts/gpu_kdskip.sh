#!/bin/bash
# scene load through the unmodified CLI with and without the CPU kd-tree build (config-2 mesh, 1.3 M triangles)
mkdir -p gpurun_out
python -m pytest tests/test_overlay_gpu.py -m gpu -q -k kdtree -s 2>&1 | tail -25 > gpurun_out/kdskip_test.log
cat gpurun_out/kdskip_test.log
python - <<'PY' > gpurun_out/kdskip_load.log 2>&1
import os, sys, subprocess, time, re
ROOT = os.getcwd(); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as g
import oracle as O
S = g.load_package().scenes
import numpy as np
td = "/tmp/kdskip"; os.makedirs(td, exist_ok=True)
sd = S.SceneDesc()
V, N, F = S.displaced_sphere(8, seed=1)
sd.add_mesh(V, N, F, bsdf=dict(type="diffuse"))
sd.add_constant_env((0.5, 0.5, 0.5))
sd.set_sensor(S.look_at((0, 0, -3.0), (0, 0, 0), (0, 1, 0)), 45.0, 256, 256, rfilter="box", spp=4, sampler="independent")
sd.set_integrator("path", max_depth=5)
xml = sd.to_xml(td, "c2mesh")
PRE = os.path.join(ROOT, "mitsuba-im_b200/plugin/overlay/lib/libmtsb_overlay.so"); OV = os.path.join(ROOT, "mitsuba-im_b200/plugin/overlay")
def run(tag, env_extra, overlay):
    env = dict(os.environ, **env_extra)
    cmd = [os.path.join(O.REF_DIR, "mitsuba"), "-C", "-p", str(os.cpu_count())] + (["-a", OV] if overlay else []) + ["-o", os.path.join(td, tag + ".pfm"), xml]
    t0 = time.perf_counter(); r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env); wall = time.perf_counter() - t0
    kd = re.findall(r"Finished -- took (\d+) ms", r.stdout); rt = re.search(r"Render time: ([0-9.]+)(ms|s|m)", r.stdout)
    print("%-34s rc=%d process wall %.2f s, render %s, kd-tree build %s ms, skipped=%s" % (tag, r.returncode, wall, rt.group(0) if rt else None, max([int(x) for x in kd]) if kd else None, "not built" in r.stdout))
run("stock", {}, False)
run("gpupath", {}, True)
run("gpupath+preload(kd build kept)", {"LD_PRELOAD": PRE, "MTSB_OVERLAY_KDBUILD": "1"}, True)
run("gpupath+preload(no kd build)", {"LD_PRELOAD": PRE}, True)
PY
cat gpurun_out/kdskip_load.log
