set -x
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --impl reference --no-cut > gpurun_out/bench_s10_reference_n2.json 2> gpurun_out/bench_s10_reference_n2.err; echo "ref exit $?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-cut > gpurun_out/bench_s10_ours_n2.json 2> gpurun_out/bench_s10_ours_n2.err; echo "ours exit $?"
tail -n 3 gpurun_out/bench_s10_ours_n2.err
python - <<'PY'
import json
for f in ("gpurun_out/bench_s10_reference_n2.json","gpurun_out/bench_s10_ours_n2.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d.get("n_gpus"), d.get("value"), d.get("ms_per_step"), d.get("e2e"), d.get("scaling"), sep="\n  ")
    except Exception as e: print(f, "ERR", e)
PY
