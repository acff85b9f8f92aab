# cost-driven collapse: sweep of the leaf cost per primitive (relative to a node) on config 2 and config 3
for CP in 0.4 0.55 0.7 0.85 1.0 1.3; do
  MTSB_DP_CPRIM=$CP python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/dp_$CP.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/dp_$CP.json"))
print("cprim=$CP", round(d["value"]), "Mrays/s  coherent", round(d["detail"]["coherent_closest_mrays"]), "shadow", round(d["detail"]["shadow_any_mrays"]), "any", round(d["detail"]["random_any_mrays"]), "nodes/ray %.2f tris/ray %.2f"%(d["roofline"]["nodes_per_ray"], d["roofline"]["prim_tests_per_ray"]))
PY
  MTSB_DP_CPRIM=$CP python bench_render.py --config 3 --no-cpu --spp3 32 > gpurun_out/dp_c3_$CP.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/dp_c3_$CP.json')); v=list(d.values())[0]; print('cprim=$CP config3', round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s')"
done
