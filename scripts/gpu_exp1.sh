# r02b: device-driven bounce graph — tests, probe of the conditional node, parity dump, render timings graph vs plain
mkdir -p gpurun_out
./scripts/exp/condgraph_probe; echo "probe rc=$?"
timeout 1200 python -m pytest tests -m gpu -q -rA > gpurun_out/r02b_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed|^FAILED|^ERROR" gpurun_out/r02b_pytest.log | tail -20
timeout 600 python scripts/exp/parity_dump.py 2>&1 | tail -8
for env in "MTSB_GRAPH=1" "MTSB_GRAPH=0" "MTSB_TAIL_PATHS=0" "MTSB_TAIL_PATHS=32768" "MTSB_TAIL_PATHS=524288"; do
  echo "== $env"; env $env timeout 600 python bench_render.py --config 1 3 5 --no-cpu --spp3 64 > gpurun_out/r02b_render_$env.json 2> gpurun_out/r02b_render_$env.err || tail -5 gpurun_out/r02b_render_$env.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r02b_render_$env.json"))
for k,v in d.items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ("samples_per_s","ms","fps","time_to_first_frame_ms","launches","launches_last_frame","e2e_ms")})
PY
done
