# sweep of the triangle-phase knob on config 2 (short runs) + new parity tests
python -m pytest tests/test_multigpu_gpu.py -x -q 2>&1 | tail -3
for T in 4 6 8 12; do
  MTSB_TRI_MIN_LANES=$T python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/sweep_$T.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/sweep_$T.json"))
print("T=$T", round(d["value"]), "Mrays/s  coherent", round(d["detail"]["coherent_closest_mrays"]), "shadow", round(d["detail"]["shadow_any_mrays"]), "any", round(d["detail"]["random_any_mrays"]), d["clocks"])
PY
  MTSB_TRI_MIN_LANES=$T python bench_render.py --config 3 --no-cpu --spp3 64 > gpurun_out/sweep_c3_$T.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/sweep_c3_$T.json')); v=list(d.values())[0]; print('T=$T config3', round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s')"
done
