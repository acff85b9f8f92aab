python -m pytest tests/test_plugin_gpu.py -x -q 2>&1 | tail -15
