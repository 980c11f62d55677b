set -x
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference > gpurun_out/bench_s12_reference.json 2> gpurun_out/bench_s12_reference.err; echo "ref exit $?"
timeout 900 python bench.py > gpurun_out/bench_s12_ours.json 2> gpurun_out/bench_s12_ours.err; echo "ours exit $?"
tail -n 3 gpurun_out/bench_s12_ours.err
python - <<'PY'
import json
for f in ("gpurun_out/bench_s12_reference.json","gpurun_out/bench_s12_ours.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d.get("value"), d.get("ms_per_step"), d.get("e2e"), d.get("roofline"), d.get("cpu_baseline"), d.get("clocks"), (d.get("train") or {}), sep="\n  ")
    except Exception as e: print(f, "ERR", e)
PY
bash tools/job29.sh
