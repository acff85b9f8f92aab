for R in 8 16 32 64 100; do
  rm -f mitsuba-im_b200/lib/obj/bvh.o
  make -C mitsuba-im_b200/csrc -j4 EXTRA=-DPLOC_RADIUS=$R > /dev/null 2>&1 || echo build failed
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/ploc_$R.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/ploc_$R.json"))
b=d["detail"]["build"]
print("R=$R", round(d["value"]), "Mrays/s  coherent", round(d["detail"]["coherent_closest_mrays"]), "shadow", round(d["detail"]["shadow_any_mrays"]), "nodes/ray %.2f tris/ray %.2f"%(d["roofline"]["nodes_per_ray"], d["roofline"]["prim_tests_per_ray"]), "build %.1f ms sah %.2f nodes %d"%(b["build_ms"], b["sah_cost"], b["nodes"]))
PY
done
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path --builder lbvh > gpurun_out/ploc_lbvh.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/ploc_lbvh.json')); b=d['detail']['build']; print('LBVH', round(d['value']), 'nodes/ray %.2f'%d['roofline']['nodes_per_ray'], 'build %.1f ms sah %.2f'%(b['build_ms'], b['sah_cost']))"
