for rep in 1 2; do
for F in 4 8; do for M in 0 1; do
  MTSB_FETCH_MIN_IDLE=$F MTSB_SORT_MATERIALS=$M python bench_render.py --config 1 --no-cpu > gpurun_out/ab_${F}_$M.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/ab_${F}_$M.json'))
for k,v in d.items(): print('F=$F sort=$M', k, round(v['samples_per_s']/1e6,1),'Msamples/s', v['launches'])"
done; done; done
