set -x
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --no-cut --no-cpu-baseline > gpurun_out/bench_s11_ours_n$N.json 2> gpurun_out/bench_s11_ours_n$N.err; echo "ours exit $?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --impl reference --no-cut > gpurun_out/bench_s11_reference_n$N.json 2> gpurun_out/bench_s11_reference_n$N.err; echo "ref exit $?"
tail -n 3 gpurun_out/bench_s11_ours_n$N.err
python - $N <<'PY'
import json, sys
N=sys.argv[1]
for f in ("gpurun_out/bench_s11_ours_n%s.json"%N,"gpurun_out/bench_s11_reference_n%s.json"%N):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d.get("n_gpus"), d.get("value"), d.get("ms_per_step"), d.get("e2e"), d.get("clocks"), sep="\n  ")
    except Exception as e: print(f, "ERR", e)
PY
