set -x
nvidia-smi --query-gpu=name,memory.total --format=csv
nproc
python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r1_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r1_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo "bench rc=$?"
cat gpurun_out/r1_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err; echo "ref rc=$?"
cat gpurun_out/r1_bench_ref.json
