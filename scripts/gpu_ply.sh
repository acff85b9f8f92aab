# device PLY decode — parity tests, then the load comparison on the config-3 meshes (bench_render.py --config ply)
python -m pytest tests/test_ply_gpu.py -m gpu -q -x 2>&1 | tail -3
python bench_render.py --config ply > gpurun_out/ply_bench.json 2> gpurun_out/ply_bench.err; tail -3 gpurun_out/ply_bench.err; cat gpurun_out/ply_bench.json
