set -x
mkdir -p gpurun_out
timeout 300 python tools/bench_eval_path.py --k 24 --frames 8 --no-reference > gpurun_out/eval_plain.json 2> gpurun_out/eval_plain.err; echo "plain exit $?"; cat gpurun_out/eval_plain.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"lodgs" -c 400 --csv --log-file gpurun_out/launches_eval_s15.csv python tools/bench_eval_path.py --k 24 --frames 8 --no-reference > gpurun_out/ncu_l.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"cut_flat|cut_weights|preprocess_fwd|blend_fwd|lod_patch" --launch-skip 15 -c 5 -o gpurun_out/full_eval_s15 -f python tools/bench_eval_path.py --k 24 --frames 8 --no-reference > gpurun_out/ncu_full_eval.log 2>&1; echo "ncu full exit $?"
ls -la gpurun_out/*.ncu-rep; tail -3 gpurun_out/ncu_full_eval.log
