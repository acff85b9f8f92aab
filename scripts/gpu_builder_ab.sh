make -C mitsuba-im_b200/csrc -j4 > /dev/null 2>&1
for B in lbvh ploc; do
  python bench_render.py --config 1 3 --no-cpu --spp3 64 --builder $B > gpurun_out/bld_$B.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/bld_$B.json'))
for k,v in d.items(): print('$B', k, round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s', 'build %.1f ms nodes %d sah %.2f'%(v['build']['build_ms'], v['build']['nodes'], v['build']['sah_cost']))"
done
