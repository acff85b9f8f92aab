python -m pytest tests -m gpu -q 2>&1 | tail -8
for i in 1 2 3 4 5 6; do python -m pytest tests/test_plugin_gpu.py -q -k gaussian 2>&1 | grep -E "passed|failed|assert np" | head -3; done
