# bench.py under torchrun at N GPUs of one box (N = $1), the way the driver launches it
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02s_bench_n$N.json 2> gpurun_out/r02s_bench_n$N.err; echo "bench rc=$?"; tail -2 gpurun_out/r02s_bench_n$N.err
python - <<PY
import json
d=[json.loads(l) for l in open("gpurun_out/r02s_bench_n$N.json") if l.startswith("{")][0]
print("N=$N value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ceiling", round(d["e2e"]["copy_only_ceiling"]), d["clocks"])
print({k: d.get(k) for k in ("config4",)})
PY
