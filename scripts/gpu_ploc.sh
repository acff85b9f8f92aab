# device-driven PLOC rounds / collapse levels: tests that exercise both builders, then build times on configs 1, 2 (via bench.py) and 3
python -m pytest tests/test_trace_gpu.py tests/test_multigpu_gpu.py -m gpu -q 2>&1 | tail -3
for B in ploc lbvh; do
  python bench_render.py --config 1 3 --no-cpu --spp3 16 --builder $B > gpurun_out/bld_$B.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/bld_$B.json'))
for k,v in d.items(): print('$B', k[:8], round(v['samples_per_s']/1e6,1),'Msamples/s', 'build %.2f ms nodes %d depth %s sah %.3f'%(v['build']['build_ms'], v['build']['nodes'], v['build'].get('depth'), v['build'].get('sah_cost', 0)))"
done
python - <<'PY'
import sys, importlib
sys.path.insert(0, '.')
B = importlib.import_module('mitsuba-im_b200.binding'); scenes = importlib.import_module('mitsuba-im_b200.scenes')
import bench; sd = bench.build_mesh(scenes)     # config 2: 1 310 720 triangles
ctx = B.Context(0)
for name, builder in (('ploc', B.BUILDER_PLOC), ('lbvh', B.BUILDER_LBVH)):
    for rep in range(3):
        sc = sd.upload(ctx, builder)
        st = sc.build_stats(); print(name, 'config2 build %.2f ms nodes %d depth %s' % (st['build_ms'], st['nodes'], st.get('depth')))
        del sc
PY
