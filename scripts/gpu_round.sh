# one GPU round: parity tests, smoke, bench (both arms). usage: bash scripts/gpu_round.sh <tag>
TAG=${1:-rXX}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; nproc
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference arm rc=$?"; tail -c 600 gpurun_out/${TAG}_bench_reference.json
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
print({k:d[k] for k in ("value","ms_per_step")}, d["e2e"]["value"], d["roofline"]["frac"], d["clocks"])
for k,v in d["detail"].items():
    if k!="path": print(k, v if not isinstance(v,dict) else "")
for k,v in d["detail"].get("path",{}).items():
    print(k, {a:b for a,b in v.items() if a in ("samples_per_s","ms","fps","time_to_first_frame_ms","latency_ms_p99","total_ms","render_ms","reduce_ms","mrays_per_s","avg_path_length","film_weight_per_pixel")})
PY
