for T in 1 2 4 8 16 32; do
  MTSB_FETCH_MIN_IDLE=$T python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/sweep_$T.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/sweep_$T.json"))
print("F=$T", round(d["value"]), "Mrays/s  coherent", round(d["detail"]["coherent_closest_mrays"]), "shadow", round(d["detail"]["shadow_any_mrays"]), "any", round(d["detail"]["random_any_mrays"]))
PY
done
for T in 2 4 8 16; do
  MTSB_FETCH_MIN_IDLE=$T python bench_render.py --config 3 --no-cpu --spp3 64 > gpurun_out/sweep_c3_$T.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/sweep_c3_$T.json')); v=list(d.values())[0]; print('F=$T config3', round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s')"
done
