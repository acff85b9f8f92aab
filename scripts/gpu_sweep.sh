# sweep of the triangle-phase knob on config 2 (short runs) + parity tests
python -m pytest tests/test_trace_gpu.py tests/test_render_gpu.py tests/test_multigpu_gpu.py -x -q 2>&1 | tail -4
for T in 1 4 8 12 16 24 33; do
  MTSB_TRI_MIN_LANES=$T python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/sweep_$T.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/sweep_$T.json"))
print("T=$T", round(d["value"]), "Mrays/s  coherent", round(d["detail"]["coherent_closest_mrays"]), "shadow", round(d["detail"]["shadow_any_mrays"]), "any", round(d["detail"]["random_any_mrays"]), "nodes/ray %.2f tris/ray %.2f"%(d["roofline"]["nodes_per_ray"], d["roofline"]["prim_tests_per_ray"]))
PY
done
for T in 1 12 33; do
  MTSB_TRI_MIN_LANES=$T python bench_render.py --config 3 --no-cpu > gpurun_out/sweep_c3_$T.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/sweep_c3_$T.json')); v=list(d.values())[0]; print('T=$T config3', round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s')"
done
