#!/usr/bin/env python
"""Turns the ncu output of scripts/gpu_ncu.sh (gpurun_out/) into the tracked summaries under profiles/.
usage: python scripts/summarise_ncu.py r01"""
import collections
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__cycles_active.avg",
        "sm__inst_executed_pipe_lsu.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum"]


def kernel_source_sha1():
    """Hash of the sources the traversal kernel and its tree are compiled from; bench.py prints roofline.traffic only
    while profiles/traffic.json carries the hash of the sources it runs (a stale capture reads as null)."""
    import hashlib
    h = hashlib.sha1()
    for f in ("trace.cu", "trace.cuh", "bvh.cu", "common.cuh"):
        h.update(open(os.path.join(ROOT, "mitsuba-im_b200", "csrc", f), "rb").read())
    return h.hexdigest()


def launches(name, tag):
    path = os.path.join(OUT, name)
    if not os.path.exists(path):
        return
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in csv.DictReader(io.StringIO("".join(lines))):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        u = r["Metric Unit"]
        v = v / 1e3 if u.startswith("n") else (v * 1e3 if u.startswith("m") else v)
        k = r["Kernel Name"].split("(")[0]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(PROF, "%s_launches_%s.csv" % (tag, name.split("_")[1].split(".")[0])), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)\n")
        f.write("kernel,launches,total_us,share_pct\n")
        for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write("%s,%d,%.1f,%.2f\n" % (k, v[0], v[1], 100 * v[1] / tot))


def full(rep, tag, what):
    path = os.path.join(OUT, rep)
    if not os.path.exists(path):
        return {}
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")].split("(")[0]}
        for k in KEYS:
            if k in hdr:
                d[k] = r[hdr.index(k)] + " " + units[hdr.index(k)]
        for i, h in enumerate(hdr):
            if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio"):
                d[h.replace("smsp__average_warps_issue_stalled_", "stall_").replace("_per_issue_active.ratio", "")] = r[i]
        out.append(d)
    with open(os.path.join(PROF, "%s_ncu_full_%s.json" % (tag, what)), "w") as f:
        json.dump(out, f, indent=1)
    return out


if __name__ == "__main__":
    tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
    os.makedirs(PROF, exist_ok=True)
    launches("launches_trace.csv", tag)
    launches("launches_render.csv", tag)
    launches("launches_c3.csv", tag)
    t = full("prof_trace.ncu-rep", tag, "trace")
    full("prof_render.ncu-rep", tag, "render")
    if t:
        def num(s): return float(s.split()[0].replace(",", ""))
        def to_bytes(s):
            v, u = s.split()[0], s.split()[1]
            return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        d = t[0]
        json.dump({"trace_closest_bytes_per_launch": to_bytes(d["dram__bytes_read.sum"]) + to_bytes(d["dram__bytes_write.sum"]),
                   "trace_closest_warp_inst_per_launch": num(d["smsp__inst_executed.sum"]),
                   "kernel_source_sha1": kernel_source_sha1(),
                   "source": "%s_ncu_full_trace.json (ncu --set full, 16 777 216 chord rays, 1.31 M triangles)" % tag},
                  open(os.path.join(PROF, "traffic.json"), "w"), indent=1)
