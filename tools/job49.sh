mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-cut > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "ours exit $?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_ours.json','gpurun_out/bench_ref.json'):
    try:
        d=json.load(open(f)); print(f, d['value'], d['ms_per_step'], d['e2e']['value'], d.get('train'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -n 3 gpurun_out/bench_ours.err; tail -n 3 gpurun_out/bench_ref.err
