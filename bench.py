#!/usr/bin/env python
"""bench.py — headline benchmark: BASELINE.json config 2, closest-hit Mrays/s on a 1.3 M-triangle synthetic mesh.

One "step" = one closest-hit pass of libmtsb200 over a batch of 16 777 216 incoherent chord rays (the distribution of
the reference's own `mtsutil kdbench`, src/utils/kdbench.cpp:226-230).  `value` is timed on the device with the rays
already resident in HBM; `e2e` is the same query through the host-pointer C-ABI call (pinned host buffers, H2D of the
rays and D2H of the hit records inside the timed region).  Coherent primary rays, shadow (any-hit) rays and the
wavefront path tracer are reported in `detail` on the same JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N > 1 is launched under torchrun (one rank per GPU); rays are sharded across ranks with no data-path collective
(weak scaling: every rank traces its own 16 M-ray batch against a replicated scene).  detail.path also carries config 4
(3840x2160, 1024 spp) split by sample index over the N ranks with one NCCL film reduce at the end (strong scaling).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

SUBDIV = 8                      # 20 * 4^8 = 1 310 720 triangles
N_RAYS = 1 << 24                # 16 777 216
BYTES_PER_RAY_CLOSEST = 48      # 32 B ray in + 16 B hit out
BYTES_PER_RAY_ANY = 36          # 32 B in + 4 B out (the device writes 1 B)
CPU_SAMPLE_RAYS = 1 << 21       # bounded sample for the CPU legs


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.rows, self.proc, self.device = [], None, device

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def build_mesh(S):
    V, N, F = S.displaced_sphere(SUBDIV, seed=1)
    sd = S.SceneDesc()
    sd.add_mesh(V, N, F)
    sd.add_point_light((0, 5, 0), (1, 1, 1))
    return sd


def reference_arm(args, rank, world):
    """The reference's own CPU implementation (oracle/_ref: unmodified ShapeKDTree::rayIntersect via oracle/raydump.cpp)
    on all host threads, on a bounded sample of the same workload."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    pkg = graft.load_package()
    S = pkg.scenes
    import tempfile
    cores = os.cpu_count() or 1
    sd = build_mesh(S)
    rays = S.chord_rays(CPU_SAMPLE_RAYS, radius=1.1, seed=42)
    with tempfile.TemporaryDirectory() as td:
        xml = sd.to_xml(td, "bench")
        if O.have_reference():
            _, tm = O.ref_trace(xml, rays, threads=cores, repeats=args.warmup + args.steps)
            times = tm.get("closest_ms_all", [tm["closest_ms"]])[args.warmup:] or [tm["closest_ms"]]
            ms = float(np.mean(times))
            kind, build_ms = "reference", tm.get("build_ms")
        else:
            osc = O.OracleScene(sd)
            sub = rays[:4096]
            t0 = time.perf_counter()
            osc.closest(sub)
            ms = (time.perf_counter() - t0) * 1e3 * (rays.shape[0] / sub.shape[0])
            kind, build_ms, cores = "port", None, 1
    value = rays.shape[0] / ms / 1e3
    sample = "%d of the %d chord rays per step, %d threads, kd-tree build (%s ms) excluded" % (rays.shape[0], N_RAYS, cores, build_ms)
    line = {"impl": "reference", "metric": "closest-hit Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "config2: closest-hit, 1310720-triangle displaced icosphere, incoherent chord rays"},
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def hit_parity(rays, ref, hits):
    """Hit records of the reference's kd-tree (oracle/raydump.cpp: t,u,v at columns 0-2, global primitive at 4) against
    the device's on the same rays: share of rays with the same primitive or a tie at equal t (north_star: >= 99.99 %),
    and the relative t / barycentric differences on the rays that agree."""
    r = np.asarray(rays).view(np.float32).reshape(-1, 8)
    ref_t, t = ref[:, 0], hits["t"]
    ref_hit, hit = np.isfinite(ref_t), np.isfinite(t)
    ref_prim = np.where(ref_hit, ref[:, 4].copy().view(np.int32), -1).astype(np.int64)
    prim = np.where(hit, hits["prim"].view(np.int32), -1).astype(np.int64)
    scale = np.maximum(np.maximum(np.abs(np.where(ref_hit, ref_t, 0)), np.abs(r[:, 0:3]).max(axis=1)), 1.0)
    with np.errstate(invalid="ignore"):
        tie = ref_hit & hit & (np.abs(t - ref_t) <= 1e-5 * scale)
    same = prim == ref_prim
    m = same & ref_hit
    dt = np.abs(t[m] - ref_t[m]) / scale[m]
    return {"n": int(len(r)), "agree": float((same | tie).mean()), "same_primitive": float(same.mean()),
            "hit_miss_agree": float((hit == ref_hit).mean()), "max_dt": float(dt.max()), "p9999_dt": float(np.quantile(dt, 0.9999)),
            "share_dt_over_1e-5": float((dt > 1e-5).mean()),
            "max_du": float(np.abs(hits["u"][m] - ref[m, 1]).max()), "max_dv": float(np.abs(hits["v"][m] - ref[m, 2]).max()),
            "vs": "oracle/_ref ShapeKDTree::rayIntersect on the first %d chord rays of the timed batch; dt relative to max(t, |o|, 1), unweighted" % len(r)}


def cpu_baseline(S, sd, cores, sc=None):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    import tempfile
    rays = S.chord_rays(CPU_SAMPLE_RAYS, radius=1.1, seed=42)
    if not O.have_reference():
        osc = O.OracleScene(sd)
        sub = rays[:2048]
        t0 = time.perf_counter()
        osc.closest(sub)
        dt = time.perf_counter() - t0
        return {"value": sub.shape[0] / dt / 1e6, "unit": "Mrays/s", "cores": 1, "kind": "port",
                "sample": "2048 rays, brute-force C restatement (no kd-tree)"}
    with tempfile.TemporaryDirectory() as td:
        xml = sd.to_xml(td, "bench")
        ref, tm = O.ref_trace(xml, rays, threads=cores, repeats=2)
    parity = hit_parity(rays, ref, sc.ray_intersect(rays)) if sc is not None else None
    return {"value": tm["closest_mrays"], "parity": parity, "unit": "Mrays/s", "cores": cores, "kind": "reference",
            "sample": "%d of the %d chord rays, best of 2, kd-tree build %s ms excluded; any-hit %.2f Mrays/s" % (
                rays.shape[0], N_RAYS, tm.get("build_ms"), tm.get("any_mrays", 0.0)),
            "kd_build_ms": tm.get("build_ms")}


_REAL_STDOUT = None


def quiet_stdout():
    """Routes fd 1 to stderr for the rest of the run so that library chatter (NCCL prints its version banner to stdout
    when NCCL_DEBUG=VERSION) cannot land beside the ONE JSON line, which emit() writes to the original stdout."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-path", action="store_true", help="skip the path-tracer detail runs (short runs under ncu)")
    ap.add_argument("--no-plugin", action="store_true", help="skip the renders through the reference CLI + overlay plugin (detail.path.plugin)")
    ap.add_argument("--no-config4", action="store_true", help="skip the 3840x2160 multi-GPU render (config 4)")
    ap.add_argument("--spp4", type=int, default=1024, help="samples per pixel of config 4 (BASELINE: 1024)")
    ap.add_argument("--builder", default="lbvh", choices=["lbvh", "ploc"],
                    help="BVH builder of the config-2 mesh (measured: LBVH 2.44, PLOC 2.08 Grays/s on the chord rays; PLOC wins on configs 1/3)")
    ap.add_argument("--rays", type=int, default=N_RAYS)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    pkg = graft.load_package()
    S, B = pkg.scenes, pkg.binding
    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = B.Context(local)
    torch.cuda.set_device(local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local))

    sd = build_mesh(S)
    builder = B.BUILDER_PLOC if args.builder == "ploc" else B.BUILDER_LBVH
    sc = sd.upload(ctx, builder)
    bstats = sc.build_stats()
    n = args.rays
    rays_h = S.chord_rays(n, radius=1.1, seed=42 + rank)
    rays_d = torch.from_numpy(rays_h.view(np.float32).reshape(-1, 8)).cuda()
    hits_d = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()

    def step():
        sc.trace_device("closest", rays_d.data_ptr(), n, hits_d.data_ptr())

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # the clock sampler starts before the warm-up (nvidia-smi needs a few hundred ms to deliver its first row, longer on
    # an 8-GPU box) and the warm-up keeps the same kernel running until the first rows are in, so that the rows recorded
    # up to the end of the timed region are all taken under this load
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        step()
    t_w = time.perf_counter()
    while len(sampler.rows) < 2 and time.perf_counter() - t_w < 3.0 and sampler.proc is not None:
        step()
        torch.cuda.synchronize()
    barrier()
    rows_before = len(sampler.rows)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kern_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with torch.cuda.stream(stream):
        e0.record(stream)
        for k in range(args.steps):
            kern_ev[k][0].record(stream)
            step()
            kern_ev[k][1].record(stream)
        e1.record(stream)
    barrier()
    # keep the load on until at least two rows fall inside/after the timed region (short runs finish between two samples)
    t_w = time.perf_counter()
    while len(sampler.rows) < rows_before + 2 and time.perf_counter() - t_w < 1.0 and sampler.proc is not None:
        step()
        torch.cuda.synchronize()
    clocks = sampler.stop()
    ms_total = e0.elapsed_time(e1)
    if world > 1:
        tt = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
    ms_step = ms_total / args.steps
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in kern_ev]))
    value = world * n / ms_step / 1e3

    # hit rate / traversal statistics (outside the timed region)
    hits_h = hits_d.cpu().numpy()
    hit_rate = float(np.isfinite(hits_h[:, 0]).mean())
    cnt = sc.trace_count_device(rays_d.data_ptr(), min(n, 1 << 22))
    nodes_per_ray, prims_per_ray = cnt["nodes"] / max(cnt["rays"], 1), cnt["prims"] / max(cnt["rays"], 1)

    # --- end-to-end through the host-pointer C ABI (pinned buffers; H2D + D2H inside the timed region) ---
    pin_r = B.PinnedArray((n,), B.ray_dtype)
    pin_h = B.PinnedArray((n,), B.hit_dtype)
    pin_r.array[:] = rays_h
    for _ in range(2):
        sc.ray_intersect(pin_r.array, out=pin_h.array)
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(3, args.steps // 2)
    for _ in range(e2e_steps):
        sc.ray_intersect(pin_r.array, out=pin_h.array)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    if world > 1:
        tt = torch.tensor([e2e_ms], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms = float(tt.item())
    assert np.array_equal(pin_h.array.view(np.uint32).reshape(-1, 4), hits_h.view(np.uint32)), "host-path hits differ from device-path hits"
    e2e_value = world * n / e2e_ms / 1e3
    # the ceiling the host-buffer path runs against, measured in the same processes: the step's H2D (rays) and D2H (hits)
    # alone, concurrently on two streams, all ranks at once — no kernel.  With N ranks this is the N-process aggregate
    # of the box's host memory system / PCIe root, which is what stops e2e from scaling.
    t_in = torch.from_numpy(pin_r.array.view(np.uint8).reshape(-1))
    t_out = torch.from_numpy(pin_h.array.view(np.uint8).reshape(-1))
    d_in = torch.empty(t_in.shape[0], dtype=torch.uint8, device="cuda")
    d_out = torch.empty(t_out.shape[0], dtype=torch.uint8, device="cuda")
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    copy_ms = []
    for _ in range(4):
        barrier()
        t0 = time.perf_counter()
        with torch.cuda.stream(s_in):
            d_in.copy_(t_in, non_blocking=True)
        with torch.cuda.stream(s_out):
            t_out.copy_(d_out, non_blocking=True)
        s_in.synchronize()
        s_out.synchronize()
        copy_ms.append((time.perf_counter() - t0) * 1e3)
    copy_ms = min(copy_ms[1:])
    if world > 1:
        tt = torch.tensor([copy_ms], device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        copy_ms = float(tt.item())
    copy_ceiling = world * n / copy_ms / 1e3
    sc.ray_intersect(pin_r.array, out=pin_h.array)      # the copy test overwrote the pinned hit buffer
    del d_in, d_out

    # --- other ray sets of config 2 (detail only) ---
    detail = {}
    if rank == 0:
        # the other builder on the same rays (both are first-class in the C ABI: mtsb_scene_commit(scene, builder))
        other = B.BUILDER_LBVH if builder == B.BUILDER_PLOC else B.BUILDER_PLOC
        sc2 = sd.upload(ctx, other)
        for _ in range(2):
            sc2.trace_device("closest", rays_d.data_ptr(), n, hits_d.data_ptr())
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            a.record(stream)
            for _ in range(3):
                sc2.trace_device("closest", rays_d.data_ptr(), n, hits_d.data_ptr())
            b.record(stream)
        torch.cuda.synchronize()
        detail["closest_mrays_%s" % ("lbvh" if other == B.BUILDER_LBVH else "ploc")] = n / (a.elapsed_time(b) / 3) / 1e3
        detail["build_other"] = sc2.build_stats()
        sc2.close()
        # the build once more: 3 of 12 recorded runs read 0.7-0.8 s for the build above (14 ms everywhere else; the phases add up to
        # 14 ms under a per-phase trace, profiles/r02r_build_trace.log) — a host-side stall inside the event window, right after
        # the clock sampler's nvidia-smi subprocess is torn down.  Both readings are reported.
        sc3 = sd.upload(ctx, other)
        detail["build_other"]["build_ms_again"] = sc3.build_stats()["build_ms"]
        sc3.close()
        sc.trace_device("closest", rays_d.data_ptr(), n, hits_d.data_ptr())      # hits_d back to the headline tree's answers
        torch.cuda.synchronize()
        res = 4096 if n >= (1 << 24) else int(np.sqrt(n))
        coh = S.pinhole_rays(res, origin=(0, 0.3, -3.0))
        coh_d = torch.from_numpy(coh.view(np.float32).reshape(-1, 8)).cuda()
        surf = sc.ray_intersect_surface(coh[:: max(1, coh.shape[0] // (1 << 20))])
        pts = surf["p"][np.isfinite(surf["t"])]
        sh = S.shadow_rays(np.tile(pts, (int(np.ceil(coh.shape[0] / max(len(pts), 1))), 1))[: coh.shape[0]], light_center=(0, 3.0, 0.5))
        sh_d = torch.from_numpy(sh.view(np.float32).reshape(-1, 8)).cuda()
        out_any = torch.empty(sh.shape[0], dtype=torch.uint8, device="cuda")
        # the like-for-like figure for ShapeKDTree::rayIntersect(ray, its): the full intersection record (92 B) per ray
        surf_d = torch.empty((n, 23), dtype=torch.float32, device="cuda")
        for name, mode, buf, out in (("coherent_closest", "closest", coh_d, hits_d), ("shadow_any", "any", sh_d, out_any),
                                     ("random_any", "any", rays_d, out_any), ("random_surface_record", "surface", rays_d, surf_d)):
            m = min(buf.shape[0], out.shape[0])
            for _ in range(2):
                sc.trace_device(mode, buf.data_ptr(), m, out.data_ptr())
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                a.record(stream)
                for _ in range(5):
                    sc.trace_device(mode, buf.data_ptr(), m, out.data_ptr())
                b.record(stream)
            torch.cuda.synchronize()
            detail[name + "_mrays"] = m / (a.elapsed_time(b) / 5) / 1e3
        if not args.no_path:
            import bench_render
            detail["path"] = bench_render.run(pkg, ctx, stream, configs=(1, 3, 5, "vol", "ply") + (() if args.no_plugin or world > 1 else ("plugin",)), cpu=not args.no_cpu_baseline and world == 1)

    # --- config 4: 3840x2160, 1024 spp split by sample index over the ranks (strong scaling), one NCCL film reduce ---
    c4s = None
    if not args.no_path and not args.no_config4:
        import bench_render
        comm = None
        if world > 1:
            uid = pkg.multigpu.exchange_unique_id(dist, B.comm_unique_id, rank)
            comm = B.Comm(ctx, world, rank, uid)
        c4 = bench_render.run(pkg, ctx, stream, configs=(4,), cpu=False, rank=rank, world=world, comm=comm, spp4=args.spp4)
        c4 = list(c4.values())[0]
        barrier()
        tt = torch.tensor([c4["total_ms"], c4["render_ms"], c4["reduce_ms"]], device="cuda", dtype=torch.float64)
        ss = torch.tensor([float(c4["paths"]), float(c4["rays"]), float(c4["launches"])], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(ss, op=dist.ReduceOp.SUM)
        if rank == 0:
            total_ms = float(tt[0].item())
            c4.update(total_ms_max_over_ranks=total_ms, render_ms_max_over_ranks=float(tt[1].item()), reduce_ms_max_over_ranks=float(tt[2].item()),
                      samples_per_s=c4["width"] * c4["height"] * c4["spp"] / (total_ms * 1e-3), paths_all_ranks=int(ss[0].item()),
                      mrays_per_s=float(ss[1].item()) / (total_ms * 1e-3) / 1e6, launches_all_ranks=int(ss[2].item()), scaling="strong",
                      split="sample index s mod %d, scene replicated, ncclReduce(sum) of the %d-byte film into rank 0" % (world, c4["film_bytes"]))
            detail.setdefault("path", {})["config4_3840x2160_%dspp_%dgpu" % (args.spp4, world)] = c4
            c4s = c4
        if comm is not None:
            comm.close()

    if rank == 0:
        peak, peak_kind = measured_peaks()
        scene_bytes = bstats["node_bytes"] + bstats["tri_bytes"]
        algo_bytes = n * BYTES_PER_RAY_CLOSEST + scene_bytes
        achieved = algo_bytes / (kern_ms * 1e-3) / 1e9
        traffic, issue = None, None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp) and n == N_RAYS:
            try:
                tj = json.load(open(tp))
                sys.path.insert(0, os.path.join(ROOT, "scripts"))
                import summarise_ncu
                if tj.get("kernel_source_sha1") != summarise_ncu.kernel_source_sha1():
                    raise ValueError("profiles/traffic.json was captured from other kernel sources")
                traffic = tj.get("trace_closest_bytes_per_launch")
                # the bound that actually binds this kernel: warp instructions issued (ncu smsp__inst_executed.sum of the
                # same launch on the same rays) against SMs x 4 schedulers x max SM clock
                props = torch.cuda.get_device_properties(local)
                issue_peak = props.multi_processor_count * 4 * (clocks.get("sm_max_mhz") or 1965.0) * 1e6
                inst = tj.get("trace_closest_warp_inst_per_launch")
                if inst:
                    issue = {"warp_inst_per_launch": inst, "achieved_per_s": inst / (kern_ms * 1e-3), "peak_per_s": issue_peak,
                             "frac": inst / (kern_ms * 1e-3) / issue_peak, "source": tj.get("source")}
            except Exception:
                traffic, issue = None, None
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            cpu = cpu_baseline(S, sd, os.cpu_count() or 1, sc)
        line = {
            "metric": "closest-hit Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "config2: closest-hit, 1310720-triangle displaced icosphere (BVH8, %s), %d incoherent chord rays per GPU per step" % (args.builder, n),
                       "l2": "inputs larger than L2 (%.0f MB of rays + %.0f MB of hits per step)" % (n * 32 / 1e6, n * 16 / 1e6),
                       "parallelism": "rays sharded across %d GPU(s), scene replicated, no collective" % world},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": n * 32, "d2h_bytes_per_step": n * 16,
                    "ms_per_step": e2e_ms, "api": "mtsb_trace_closest(host pointers), pinned buffers, 2M-ray chunks pipelined over 3 streams and a 4-buffer ring",
                    "copy_only_ceiling": copy_ceiling, "copy_only_ms": copy_ms, "frac_of_copy_ceiling": e2e_value / copy_ceiling,
                    "limiter": "host<->device copies: the same bytes moved with no kernel at all, by the same %d process(es) at once, run at copy_only_ceiling" % world},
            "gpu_launches": args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_kind": peak_kind, "kernel": "trace_kernel<closest>", "kernel_ms": kern_ms,
                         "algorithmic_bytes": algo_bytes, "nodes_per_ray": nodes_per_ray, "prim_tests_per_ray": prims_per_ray,
                         "traversal_bytes_per_ray": nodes_per_ray * 80 + prims_per_ray * 48 + 48, "issue": issue},
            "cpu_baseline": cpu,
            "parity": cpu.pop("parity", None) if cpu else None,
            # the multi-GPU result that has a collective in it (strong scaling of config 4, one ncclReduce of the film)
            "config4": {k: c4s.get(k) for k in ("samples_per_s", "total_ms_max_over_ranks", "render_ms_max_over_ranks", "reduce_ms_max_over_ranks",
                                               "film_bytes", "spp", "width", "height", "scaling", "split")} if c4s else None,
            "detail": dict(detail, hit_rate=hit_rate, build=bstats),
        }
        emit(line)
    pin_r.free()
    pin_h.free()
    sc.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
