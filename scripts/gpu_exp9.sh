mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_plugin_gpu.py tests/test_overlay_gpu.py -m gpu -q 2>&1 | tail -4
python bench_render.py --config plugin > gpurun_out/r02l_plugin.json 2> gpurun_out/r02l_plugin.err; python -c "
import json; d=json.load(open('gpurun_out/r02l_plugin.json'))['plugin']
for k,v in d.items(): print(k, v if not isinstance(v,dict) else {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a!='error'} , (v.get('error','')[-300:] if isinstance(v,dict) else ''))"
MTSB_PLUGIN_TIMING=1 oracle/_ref/mitsuba -p 16 -a mitsuba-im_b200/plugin/overlay -D sampler=independent -D spp=64 -D res=512 -D rfilter=box -o /tmp/r.pfm tests/golden/cbox_config1.xml 2>&1 | grep -E "leader|responsive round" | cut -c1-260
