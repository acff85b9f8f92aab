#!/bin/bash
# volpath_simple parity on the GPU box: the fog goldens, the plugin route, then the whole render test file
mkdir -p gpurun_out
python -m pytest tests/test_render_gpu.py tests/test_plugin_gpu.py -m gpu -q -k "volpath" -s 2>&1 | tail -40 > gpurun_out/vol_tests.log
python -m pytest tests/test_render_gpu.py tests/test_plugin_gpu.py -m gpu -q -x 2>&1 | tail -15 > gpurun_out/vol_all.log
cat gpurun_out/vol_tests.log gpurun_out/vol_all.log
