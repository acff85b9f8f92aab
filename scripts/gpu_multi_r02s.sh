# 2-GPU round on the final tree: the whole GPU test suite on a 2-GPU box (the two multi-GPU tests run), then bench.py under torchrun at N=2
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 1500 python -m pytest tests -m gpu -q -rA > gpurun_out/r02s_pytest_gpu_n2.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed|^FAILED|^ERROR|skipped" gpurun_out/r02s_pytest_gpu_n2.log | tail -8
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02s_bench_n2.json 2> gpurun_out/r02s_bench_n2.err; echo "bench rc=$?"; tail -3 gpurun_out/r02s_bench_n2.err
python - <<PY
import json
d=[json.loads(l) for l in open("gpurun_out/r02s_bench_n2.json") if l.startswith("{")][0]
print("N=2 value", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["clocks"])
for k,v in d["detail"].get("path",{}).items():
    if k.startswith("config4"): print(k, {a:v.get(a) for a in ("samples_per_s","total_ms_max_over_ranks","render_ms_max_over_ranks","reduce_ms_max_over_ranks","film_weight_per_pixel","mean_radiance","split")})
print({k: d.get(k) for k in ("config4",)})
PY
