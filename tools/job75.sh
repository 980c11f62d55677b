set -x
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 )
for v in lpt tileid; do
  if [ $v = tileid ]; then export LODGS_BLEND_TILE_ID_ORDER=1; else unset LODGS_BLEND_TILE_ID_ORDER; fi
  EVAL_DEBUG=1 timeout 600 python tools/bench_eval_path.py --k 24 --frames 200 --no-reference > gpurun_out/eval_ab_$v.json 2> gpurun_out/eval_ab_$v.err; echo "eval $v exit $?"; grep "blend\|tile_order" gpurun_out/eval_ab_$v.err; cut -c1-700 gpurun_out/eval_ab_$v.json
  timeout 600 python bench.py --no-cut --no-cpu-baseline > gpurun_out/bench_ab_$v.json 2> gpurun_out/bench_ab_$v.err; echo "bench $v exit $?"; python -c "
import json;d=json.load(open('gpurun_out/bench_ab_$v.json'));print('$v',d['value'],d['ms_per_step'],d['e2e']['value'],[(k['kernel'],k['avg_ms']) for k in d['per_kernel'][:3]], d.get('train'))"
  timeout 600 python tools/bench_hier_step.py --k 21 --iters 10 > gpurun_out/hier_ab_$v.json 2> gpurun_out/hier_ab_$v.err; echo "hier $v exit $?"; python -c "
import json;d=json.load(open('gpurun_out/hier_ab_$v.json'));print('$v',{t:(e['ours'],e.get('bit_exact')) for t,e in d['taus'].items()})"
done
