mkdir -p gpurun_out
for w in 2 0 1 4 16; do
  echo "== MTSB_SHADE_WAVES=$w"; MTSB_SHADE_WAVES=$w timeout 600 python bench_render.py --config 1 3 --no-cpu --spp3 64 > gpurun_out/r02h_$w.json 2> gpurun_out/r02h_$w.err || tail -5 gpurun_out/r02h_$w.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r02h_$w.json"))
for k,v in d.items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ("samples_per_s","ms")})
PY
done
echo "== plugin"
python bench_render.py --config plugin > gpurun_out/r02h_plugin.json 2> gpurun_out/r02h_plugin.err; python -c "
import json; d=json.load(open('gpurun_out/r02h_plugin.json'))['plugin']
for k,v in d.items(): print(k, v if not isinstance(v,dict) else {a:b for a,b in v.items() if a!='error'} , (v.get('error','')[-300:] if isinstance(v,dict) else ''))"
MTSB_PLUGIN_TIMING=1 oracle/_ref/mitsuba -p 16 -a mitsuba-im_b200/plugin/overlay -D sampler=independent -D spp=64 -D res=512 -D rfilter=box -o /tmp/r.pfm tests/golden/cbox_config1.xml 2>&1 | grep -E "gpupath|SPP" | cut -c1-260
timeout 900 python -m pytest tests/test_plugin_gpu.py -m gpu -q 2>&1 | tail -5
