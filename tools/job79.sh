set -x
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -3 ) | tee gpurun_out/pytest_s16.txt
timeout 600 python bench.py --impl reference > gpurun_out/bench_s16_reference.json 2> gpurun_out/bench_s16_reference.err; echo "ref exit $?"
timeout 600 python bench.py > gpurun_out/bench_s16_ours.json 2> gpurun_out/bench_s16_ours.err; echo "ours exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/bench_s16_ours.json","gpurun_out/bench_s16_reference.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], d["e2e"], d.get("train"), d.get("gpu_launches"), d.get("clocks"))
    except Exception as e: print(f,"ERR",e)
PY
KR='regex:cut_|preprocess|blend|sort_|emit_|scan_sorted|tile_ranges|tile_order|lod_patch|publish|adam'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KR" -c 400 --csv --log-file gpurun_out/launches_s16.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cut > gpurun_out/ncu_l16.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KR" -c 400 --csv --log-file gpurun_out/launches_eval_s16.csv python tools/bench_eval_path.py --k 24 --frames 8 --no-reference > gpurun_out/ncu_le16.log 2>&1; echo "ncu eval launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"blend_fwd|blend_bwd|preprocess_fwd|preprocess_bwd|sort_onesweep|tile_order" --launch-skip 40 -c 12 -o gpurun_out/full_s16 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-cut > gpurun_out/ncu_full16.log 2>&1; echo "ncu full exit $?"
ls -la gpurun_out/*.ncu-rep gpurun_out/launches*_s16.csv
