python -m pytest tests/test_render_gpu.py tests/test_multigpu_gpu.py tests/test_plugin_gpu.py -x -q 2>&1 | tail -3
for M in 0 1; do
  MTSB_SORT_MATERIALS=$M python bench_render.py --config 1 3 --no-cpu --spp3 64 > gpurun_out/sort_$M.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/sort_$M.json'))
for k,v in d.items(): print('sort=$M', k, round(v['samples_per_s']/1e6,1),'Msamples/s', round(v['mrays_per_s']),'Mrays/s')"
done
