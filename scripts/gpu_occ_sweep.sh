for MB in 8 9 10 12; do
  rm -f mitsuba-im_b200/lib/obj/render.o mitsuba-im_b200/lib/obj/trace.o
  make -C mitsuba-im_b200/csrc -j4 EXTRA=-DMTSB_TRACE_BLOCKS_PER_SM=$MB > /dev/null 2>&1 || echo build failed
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/occ_$MB.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/occ_$MB.json"))
print("MB=$MB", round(d["value"]), "Mrays/s  coherent", round(d["detail"]["coherent_closest_mrays"]), "shadow", round(d["detail"]["shadow_any_mrays"]), "any", round(d["detail"]["random_any_mrays"]))
PY
  python bench_render.py --config 3 --no-cpu --spp3 64 > gpurun_out/occ_c3_$MB.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/occ_c3_$MB.json')); v=list(d.values())[0]; print('MB=$MB config3', round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s')"
done
