set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 500 python tools/e2e_timeline.py 2>&1 | tail -16
timeout 600 python bench.py --no-cut --no-cpu-baseline > gpurun_out/b_ours.json 2> gpurun_out/b_ours.err; echo "ours exit $?"
timeout 600 python bench.py --impl reference --no-cut > gpurun_out/b_ref.json 2> gpurun_out/b_ref.err; echo "ref exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/b_ours.json","gpurun_out/b_ref.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], d["e2e"], d.get("train"))
    except Exception as e: print(f,"ERR",e)
PY
tail -n 3 gpurun_out/b_ours.err
