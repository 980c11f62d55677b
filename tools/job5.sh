set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "^(FAILED|ERROR)|passed|failed|^E  " gpurun_out/pytest_gpu.log | tail -15 | cut -c1-300
timeout 600 python tests/gpu_diag.py > gpurun_out/parity_diag.log 2>&1; echo "diag exit $?"
grep -E "^==|FAIL|integer mismatches" gpurun_out/parity_diag.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-cut > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench exit $?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_ours.json',):
    d=json.load(open(f))
    print(d['value'], d['ms_per_step'], d['e2e'])
    for k in d.get('per_kernel',[]): print(k['kernel'], k['launches_per_step'], k['ms_per_step'], k.get('frac'))
PY
tail -3 gpurun_out/bench_ours.err
