# first GPU contact of a session: all GPU tests (no -x, to see every failure), smoke, bench. usage: bash scripts/gpu_first.sh <tag>
TAG=${1:-rXX}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; nproc
timeout 1500 python -m pytest tests -m gpu -q -rA --durations=15 > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed|PASSED|FAILED|ERROR" gpurun_out/${TAG}_pytest.log | tail -60
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
print({k:d.get(k) for k in ("value","ms_per_step","parity")}, d["e2e"]["value"], d["roofline"], d["clocks"])
for k,v in d["detail"].items():
    if k!="path": print(k, v if not isinstance(v,dict) else "")
for k,v in d["detail"].get("path",{}).items():
    print(k, {a:b for a,b in v.items() if a in ("samples_per_s","ms","fps","time_to_first_frame_ms","latency_ms_p99","total_ms","render_ms","reduce_ms","mrays_per_s","avg_path_length","film_weight_per_pixel","launches_last_frame")} if isinstance(v,dict) else v)
PY
