# builder A/B on the render configs (and PLOC search radius, a build-time knob)
for R in 16 8; do
  rm -f mitsuba-im_b200/lib/obj/bvh.o
  make -C mitsuba-im_b200/csrc -j4 EXTRA=-DPLOC_RADIUS=$R > /dev/null 2>&1 || echo build failed
  for B in ploc lbvh; do
    if [ "$R" = "8" ] && [ "$B" = "lbvh" ]; then continue; fi
    python bench_render.py --config 1 3 --no-cpu --spp3 64 --builder $B > gpurun_out/bld_$B.json 2>/dev/null
    python -c "
import json; d=json.load(open('gpurun_out/bld_$B.json'))
for k,v in d.items(): print('R=$R $B', k[:8], round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s', 'build %.1f ms nodes %d'%(v['build']['build_ms'], v['build']['nodes']))"
  done
done
