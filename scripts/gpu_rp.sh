#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_render_gpu.py tests/test_plugin_gpu.py -m gpu -q -k "roughplastic" 2>&1 | tail -40 > gpurun_out/rp_tests.log
cat gpurun_out/rp_tests.log
