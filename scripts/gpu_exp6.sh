mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rA --durations=8 > gpurun_out/r02i_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed|^FAILED|^ERROR" gpurun_out/r02i_pytest.log | tail -20
grep -E "light record|reference: .* normal|parity\[" gpurun_out/r02i_pytest.log | cut -c1-250 | head -30
bash scripts/gpu_sanitize.sh
