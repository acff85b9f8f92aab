mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv,noheader
timeout 1500 python -m pytest tests -m gpu -q -rA > gpurun_out/r02_pytest_gpu_n2.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed|^FAILED|^ERROR|skipped|plugin, classic" gpurun_out/r02_pytest_gpu_n2.log | tail -12
python bench_render.py --config plugin > gpurun_out/r02_plugin_n2.json 2> gpurun_out/r02_plugin_n2.err; python -c "
import json; d=json.load(open('gpurun_out/r02_plugin_n2.json'))['plugin']
for k,v in d.items(): print(k, v if not isinstance(v,dict) else {a:b for a,b in v.items() if a!='error'} , (v.get('error','')[-300:] if isinstance(v,dict) else ''))"
