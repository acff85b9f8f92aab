# quick check after a renderer change: render parity tests, configs 1 / 3 / 5
python -m pytest tests/test_render_gpu.py tests/test_multigpu_gpu.py -x -q 2>&1 | tail -1
python bench_render.py --config 1 3 5 --no-cpu --spp3 64 > gpurun_out/qr.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/qr.json'))
for k,v in d.items(): print(k[:7], {a:(round(b,2) if isinstance(b,float) else b) for a,b in v.items() if a in ('samples_per_s','ms','fps','time_to_first_frame_ms','latency_ms_p99','mrays_per_s','avg_path_length')})"
