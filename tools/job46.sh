mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cut_dynamic.py -m gpu -q > gpurun_out/pytest_dyn.log 2>&1; echo "pytest exit $?"
grep -E "^(FAILED|ERROR)|passed|failed|^E  " gpurun_out/pytest_dyn.log | tail -30 | cut -c1-300
