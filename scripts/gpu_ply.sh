python -m pytest tests/test_ply_gpu.py -m gpu -q -x 2>&1 | tail -15
python scripts/gpu_ply.py 6 > gpurun_out/ply_bench.json 2> gpurun_out/ply_bench.err; tail -3 gpurun_out/ply_bench.err; cat gpurun_out/ply_bench.json
