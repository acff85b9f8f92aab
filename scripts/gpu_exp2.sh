mkdir -p gpurun_out
for env in "MTSB_L2_PERSIST=0" "MTSB_L2_PERSIST=16" ; do
  echo "== $env"; env $env timeout 600 python bench_render.py --config 1 3 5 --no-cpu --spp3 64 > gpurun_out/r02c_render_$env.json 2> gpurun_out/r02c_render_$env.err || tail -5 gpurun_out/r02c_render_$env.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r02c_render_$env.json"))
for k,v in d.items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ("samples_per_s","ms","fps","time_to_first_frame_ms","launches","launches_last_frame","e2e_ms")})
PY
done
R1="python bench_render.py --config 1 --no-cpu"
MTSB_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02c_launches_render.csv $R1 > gpurun_out/ncu_render1.log 2>&1
python scripts/exp/launch_table.py gpurun_out/r02c_launches_render.csv
