mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_render_gpu.py tests/test_plugin_gpu.py tests/test_multigpu_gpu.py -m gpu -q 2>&1 | tail -6
for sa in 0 8 24 64 200; do
  echo "== MTSB_SUSPEND_AFTER=$sa"
  MTSB_SUSPEND_AFTER=$sa timeout 600 python bench_render.py --config 1 3 5 --no-cpu --spp3 64 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read())
for k,v in d.items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('samples_per_s','ms','fps','time_to_first_frame_ms','launches_last_frame','avg_path_length','mean_radiance')})"
done
