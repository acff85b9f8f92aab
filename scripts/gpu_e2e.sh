python -m pytest tests/test_trace_gpu.py -x -q 2>&1 | tail -2
for C in 2097152 1048576 524288 262144; do
  MTSB_HOST_CHUNK=$C python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/e2e_$C.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/e2e_$C.json')); print('chunk=$C value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['e2e']['ms_per_step'],2))"
done
