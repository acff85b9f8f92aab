mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rA > gpurun_out/r02k_pytest.log 2>&1; echo "pytest rc=$?"; grep -aE "passed|failed|^FAILED|^ERROR|analytic (sphere|disk)" gpurun_out/r02k_pytest.log | cut -c1-300 | tail -14
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('C2', round(d['value']), round(d['detail']['shadow_any_mrays']), round(d['detail']['coherent_closest_mrays']))"
python bench_render.py --config 1 3 5 --no-cpu --spp3 64 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read())
for k,v in d.items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('samples_per_s','ms','fps','time_to_first_frame_ms')})"
