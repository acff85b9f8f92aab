# quick check after a traversal change: parity tests, config-2 ray sets, config 3
python -m pytest tests/test_trace_gpu.py tests/test_render_gpu.py -x -q 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/quick.json 2>gpurun_out/quick.err || tail -5 gpurun_out/quick.err
python - <<PY
import json
d=json.load(open("gpurun_out/quick.json"))
print(round(d["value"]), "Mrays/s  coherent", round(d["detail"]["coherent_closest_mrays"]), "shadow", round(d["detail"]["shadow_any_mrays"]), "any", round(d["detail"]["random_any_mrays"]), "nodes/ray %.2f tris/ray %.2f"%(d["roofline"]["nodes_per_ray"], d["roofline"]["prim_tests_per_ray"]), "e2e", round(d["e2e"]["value"]))
PY
python bench_render.py --config 1 3 --no-cpu > gpurun_out/quick_c3.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/quick_c3.json'))
for k,v in d.items(): print(k, round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s')"
