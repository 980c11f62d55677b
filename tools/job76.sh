set -x
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 )
for v in cls tileid; do
  if [ $v = tileid ]; then export LODGS_BLEND_TILE_ID_ORDER=1; else unset LODGS_BLEND_TILE_ID_ORDER; fi
  EVAL_DEBUG=1 timeout 600 python tools/bench_eval_path.py --k 24 --frames 200 --no-reference > gpurun_out/eval_ab2_$v.json 2> gpurun_out/eval_ab2_$v.err; echo "eval $v exit $?"; grep "blend\|tile_order" gpurun_out/eval_ab2_$v.err; cut -c1-700 gpurun_out/eval_ab2_$v.json
  timeout 600 python bench.py --no-cut --no-cpu-baseline > gpurun_out/bench_ab2_$v.json 2> gpurun_out/bench_ab2_$v.err; echo "bench $v exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_ab2_$v.json'));print('$v',d['value'],d['ms_per_step'],d['e2e']['value'],[(k['kernel'],k['avg_ms']) for k in d['per_kernel'][:3]], d.get('train'))"
  timeout 600 python tools/bench_hier_step.py --k 21 --iters 10 > gpurun_out/hier_ab2_$v.json 2> gpurun_out/hier_ab2_$v.err; echo "hier $v exit $?"; python -c "
import json;d=json.load(open('gpurun_out/hier_ab2_$v.json'));print('$v',{t:(e['ours']['fwd_bwd_ms'],e['ours']['step_ms'],all(e.get('bit_exact').values())) for t,e in d['taus'].items()})"
done
unset LODGS_BLEND_TILE_ID_ORDER
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29532 tools/bench_dp_train.py > gpurun_out/dp_train_s15_ours_n2.json 2> gpurun_out/dp_train_s15_ours_n2.err; echo "dp ours exit $?"; tail -n 3 gpurun_out/dp_train_s15_ours_n2.err; cat gpurun_out/dp_train_s15_ours_n2.json
timeout 300 $TR --master-port 29533 tools/bench_dp_train.py --no-overlap > gpurun_out/dp_train_s15_ours_noov_n2.json 2> gpurun_out/dp_train_s15_ours_noov_n2.err; echo "dp ours exit $?"; cat gpurun_out/dp_train_s15_ours_noov_n2.json
