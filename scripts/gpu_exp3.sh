mkdir -p gpurun_out
timeout 600 python bench_render.py --config 1 3 5 --no-cpu --spp3 64 > gpurun_out/r02d_render.json 2> gpurun_out/r02d_render.err || tail -5 gpurun_out/r02d_render.err
python - <<PY
import json
d=json.load(open("gpurun_out/r02d_render.json"))
for k,v in d.items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ("samples_per_s","ms","fps","time_to_first_frame_ms","launches","launches_last_frame","e2e_ms")})
PY
cat > /tmp/c5few.py <<PY
import sys, os
sys.path.insert(0, os.getcwd())
import bench_render as br, torch
pkg = br.graft.load_package(); ctx = pkg.binding.Context(0); torch.cuda.set_device(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", 0))
S = pkg.scenes
sd = S.sphere_field(subdiv=6, nx=16, ny=8, width=1280, height=720, spp=256, max_depth=8, rfilter="gaussian")
print(br.render_config5(pkg, ctx, stream, sd, frames=6, warmup=2))
PY
for t in 131072 0; do
MTSB_GRAPH=0 MTSB_TAIL_PATHS=$t ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02d_launches_c5_tail$t.csv python /tmp/c5few.py > gpurun_out/ncu_c5_$t.log 2>&1
echo "== C5 last frame, tail=$t"
python - <<PY
import csv, io
lines=[l for l in open("gpurun_out/r02d_launches_c5_tail$t.csv") if not l.startswith("==")]
rows=[r for r in csv.DictReader(io.StringIO("".join(lines))) if r.get("Metric Name")=="gpu__time_duration.sum"]
# last frame = from the last generate_kernel on
last=max(i for i,r in enumerate(rows) if "generate_kernel" in r["Kernel Name"])
tot=0
for r in rows[last:]:
    v=float(r["Metric Value"].replace(",","")); u=r["Metric Unit"]; v = v/1e3 if u.startswith("n") else (v*1e3 if u.startswith("m") else v)
    tot+=v; print("%-28s %8.1f" % (r["Kernel Name"].split("(")[0][:28], v))
print("sum", tot)
PY
done
