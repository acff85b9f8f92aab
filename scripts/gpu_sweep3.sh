# re-sweep of the two scheduling knobs of the traversal kernel after the L2 / streaming-hint changes (config 2 + config 3)
mkdir -p gpurun_out
for kv in "8 8" "4 8" "12 8" "16 8" "8 4" "8 12" "8 16" "8 33" "16 16" "4 4"; do
  set -- $kv
  r=$(MTSB_FETCH_MIN_IDLE=$1 MTSB_TRI_MIN_LANES=$2 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-path 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['detail']['shadow_any_mrays']), round(d['detail']['coherent_closest_mrays']))")
  c3=$(MTSB_FETCH_MIN_IDLE=$1 MTSB_TRI_MIN_LANES=$2 python bench_render.py --config 3 --no-cpu --spp3 32 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(list(d.values())[0]['samples_per_s']/1e6,1))")
  echo "fetchMinIdle=$1 triMinLanes=$2 : C2 closest/shadow/coherent Mrays/s = $r ; C3 M samples/s = $c3"
done | tee gpurun_out/r02_knob_sweep.log
