python -m pytest tests/test_trace_gpu.py tests/test_render_gpu.py tests/test_multigpu_gpu.py tests/test_plugin_gpu.py -x -q 2>&1 | tail -3
for MB in 4 5 6; do
  rm -f mitsuba-im_b200/lib/obj/render.o
  make -C mitsuba-im_b200/csrc EXTRA=-DMTSB_SHADE_MIN_BLOCKS=$MB > /dev/null 2>&1 || echo build failed
  python bench_render.py --config 1 3 --no-cpu > gpurun_out/shade_$MB.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/shade_$MB.json'))
for k,v in d.items(): print('MB=$MB', k, round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s')"
done
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/quick.json 2>gpurun_out/quick.err || tail -5 gpurun_out/quick.err
python - <<PY
import json
d=json.load(open("gpurun_out/quick.json"))
print(round(d["value"]), "Mrays/s  coherent", round(d["detail"]["coherent_closest_mrays"]), "shadow", round(d["detail"]["shadow_any_mrays"]), "any", round(d["detail"]["random_any_mrays"]))
PY
