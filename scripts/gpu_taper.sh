# host-buffer ray path: tapered tail chunks (MTSB_HOST_TAPER = smallest piece, 0 = off)
python -m pytest tests/test_trace_gpu.py -x -q 2>&1 | tail -2
for T in 0 1048576 524288 262144 131072 65536; do
  MTSB_HOST_TAPER=$T python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-path > gpurun_out/taper_$T.json 2>/dev/null
  python -c "
import json; d=json.load(open('gpurun_out/taper_$T.json')); print('taper=$T value', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['e2e']['ms_per_step'],2))"
done
