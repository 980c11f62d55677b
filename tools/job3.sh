set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu.log | tail -15
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_ours.json'))
print(d['value'], d['ms_per_step'], d['e2e'])
for k in d['per_kernel']: print(k['kernel'], k['ms_per_step'], k.get('frac'))
PY
timeout 600 python tests/gpu_diag.py > gpurun_out/parity_diag.log 2>&1; echo "diag exit $?"
grep -E "^==|FAIL|integer mismatches|final_T|color " gpurun_out/parity_diag.log
