# L2 window experiments on config 2 (device-resident closest hit): hit ratio and window size of the persisting window
run() { echo "== $*"; env "$@" python bench.py --no-path --no-cpu-baseline --steps 6 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); print(round(d['value'],1), 'Mrays/s; ploc', round(d['detail']['closest_mrays_ploc'],1), 'coh', round(d['detail']['coherent_closest_mrays'],1), 'shadow', round(d['detail']['shadow_any_mrays'],1))"; }
run MTSB_NOP=1
run MTSB_L2_WINDOW_MB=15
run MTSB_L2_WINDOW_MB=40
run MTSB_L2_RATIO=0.6
run MTSB_L2_RATIO=0.8
run MTSB_L2_PERSIST=0
