set -x
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 ) | tee gpurun_out/pytest_s18.txt
python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/bench_s18_ours.json 2> gpurun_out/bench_s18_ours.err; echo "ours exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_s18_ours.json'));print(d['value'],d['ms_per_step'],d['e2e'],d.get('train'),d['gpu_launches'],d['roofline'],d['cpu_baseline'],d['clocks'])"
timeout 300 python tools/bench_eval_path.py --k 25 --frames 100 --check-every 10 > gpurun_out/eval_path_s18_k25_n1.json 2> gpurun_out/eval_path_s18_k25_n1.err; echo "k25 exit $?"; cat gpurun_out/eval_path_s18_k25_n1.json
