"""bench_render.py — path samples/s (px*spp/s) of the wavefront path tracer; imported by bench.py (detail["path"]) and
runnable on its own:  python bench_render.py [--config 1|3] [--spp N] [--cpu]

config 1: Cornell box (rectangles + cube, diffuse, area light), `path` maxDepth 5, 512x512, 64 spp
config 3: 10 485 762-triangle displaced sphere field, GGX rough copper + bk7 glass, area light + constant environment,
          `path` maxDepth 8, 1920x1080, 256 spp, Gaussian filter
Timed on the device (CUDA events on the context stream) from the first generate kernel to the last splat; the film stays
in HBM.  `e2e` adds the D2H read of the film (ImageBlock layout) — the only per-frame transfer of the plugin path.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402


def render_config(pkg, ctx, stream, sd, spp, warmup=1, builder=None):
    import torch
    B = pkg.binding
    t0 = time.perf_counter()
    sc = sd.upload(ctx, B.BUILDER_PLOC if builder is None else builder)
    upload_s = time.perf_counter() - t0
    rd = B.Render(sc, sd.camera(), sd.film(), sd.integrator_desc(), seed=1)
    w, h = sd.sensor["width"], sd.sensor["height"]
    for _ in range(warmup):
        rd.render_pass(0, min(spp, 4), spp)
    rd.sync()
    rd.clear()
    rd.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        rd.render_pass(0, spp, spp)
        e1.record(stream)
    rd.sync()
    ms = e0.elapsed_time(e1)
    st = rd.stats()
    # end to end: render + film read-back into host memory
    rd.clear()
    rd.sync()
    t0 = time.perf_counter()
    rd.render_pass(0, spp, spp)
    film = rd.film_to_host()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    b = sd.film().border
    img = film[b:film.shape[0] - b, b:film.shape[1] - b]
    out = {
        "samples_per_s": w * h * spp / (ms * 1e-3), "ms": ms, "width": w, "height": h, "spp": spp,
        "e2e_samples_per_s": w * h * spp / (e2e_ms * 1e-3), "e2e_ms": e2e_ms, "film_bytes": int(film.nbytes),
        "avg_path_length": st["depth_sum"] / max(st["paths"], 1),
        "rays_per_sample": (st["closest_rays"] + st["shadow_rays"]) / max(st["paths"], 1),
        "mrays_per_s": (st["closest_rays"] + st["shadow_rays"]) / (ms * 1e-3) / 1e6,
        "launches": st["launches"], "triangles": int(sc.primitive_count), "build": sc.build_stats(), "upload_s": upload_s,
        "mean_radiance": float(np.nanmean(img[..., 0:3] / np.maximum(img[..., 4:5], 1e-9))),
    }
    # algorithmic HBM bytes per sample: 32 B film RMW + avg path length x 444 B
    out["algorithmic_bytes_per_sample"] = 32 + out["avg_path_length"] * 444
    out["achieved_GBps"] = out["algorithmic_bytes_per_sample"] * out["samples_per_s"] / 1e9
    rd.close()
    sc.close()
    return out


def cpu_reference(pkg, sd_small, threads):
    """The reference's own renderer (oracle/_ref/mitsuba -C) on the host cores, bounded sample of config 1."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    import tempfile
    if not O.have_reference():
        return None
    with tempfile.TemporaryDirectory() as td:
        xml = sd_small.to_xml(td, "cbox")
        wall, render_s, _ = O.ref_render(xml, os.path.join(td, "o.pfm"), threads=threads, classic=True)
    n = sd_small.sensor["width"] * sd_small.sensor["height"] * sd_small.sampler["spp"]
    t = render_s if render_s else wall
    return {"samples_per_s": n / t, "cores": threads, "kind": "reference",
            "sample": "config 1 at %d spp (of 64), mitsuba -C, 'Render time' %.3f s" % (sd_small.sampler["spp"], t)}


def run(pkg, ctx, stream, configs=(1, 3), cpu=True):
    S = pkg.scenes
    out = {}
    if 1 in configs:
        out["config1_cornell_512x512_64spp"] = render_config(pkg, ctx, stream, S.cornell_box(512, 512, spp=64, max_depth=5), 64)
        if cpu:
            out["config1_cpu_reference"] = cpu_reference(pkg, S.cornell_box(512, 512, spp=16, max_depth=5), os.cpu_count() or 1)
    if 3 in configs:
        t0 = time.perf_counter()
        sd = S.sphere_field(subdiv=6, nx=16, ny=8, width=1920, height=1080, spp=256, max_depth=8, rfilter="gaussian")
        gen_s = time.perf_counter() - t0
        r = render_config(pkg, ctx, stream, sd, 256, warmup=1)
        r["scene_generation_s"] = gen_s
        out["config3_spherefield_10Mtri_1920x1080_256spp"] = r
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, nargs="*", default=[1, 3])
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import torch
    pkg = graft.load_package()
    ctx = pkg.binding.Context(0)
    torch.cuda.set_device(0)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", 0))
    print(json.dumps(run(pkg, ctx, stream, tuple(args.config), cpu=not args.no_cpu)))
