# host-buffer ray path: parity tests, then end-to-end rate three times (scripts/gpu_pcie.py gives the PCIe ceiling)
python -m pytest tests/test_trace_gpu.py tests/test_abi.py -x -q 2>&1 | tail -1
for i in 1 2 3; do
python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/pt2.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/pt2.json')); print('value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['e2e']['ms_per_step'],2))"
done
python scripts/gpu_pcie.py
