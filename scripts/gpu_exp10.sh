mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_render_gpu.py tests/test_parity_configs_gpu.py -m gpu -q 2>&1 | tail -4
for v in 1 0; do
  echo "== MTSB_SHADE_LEAN=$v"
  MTSB_SHADE_LEAN=$v timeout 600 python bench_render.py --config 1 3 5 --no-cpu --spp3 64 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read())
for k,v in d.items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('samples_per_s','ms','fps','time_to_first_frame_ms','mean_radiance')})"
done
