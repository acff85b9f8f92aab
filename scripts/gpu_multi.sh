# multi-GPU round: 2-GPU NCCL film-reduce test, then bench.py under torchrun at N = $1
N=${1:-2}
nvidia-smi --query-gpu=index,name --format=csv,noheader
python -m pytest tests/test_multigpu_gpu.py -x -q 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/multi_n$N.json 2> gpurun_out/multi_n$N.err; echo "bench rc=$?"; tail -5 gpurun_out/multi_n$N.err
python - <<PY
import json
d=[json.loads(l) for l in open("gpurun_out/multi_n$N.json") if l.startswith("{")][0]
print("N=$N value", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["clocks"])
for k,v in d["detail"].get("path",{}).items():
    if k.startswith("config4"): print(k, {a:v[a] for a in ("samples_per_s","total_ms_max_over_ranks","render_ms_max_over_ranks","reduce_ms_max_over_ranks","film_weight_per_pixel","mean_radiance","split")})
PY
