set -x
mkdir -p gpurun_out
export LODGS_LOD_CUT=1
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 )
EVAL_DEBUG=1 timeout 600 python tools/bench_eval_path.py --k 24 --frames 200 --check-every 10 > gpurun_out/eval_s17_n1.json 2> gpurun_out/eval_s17_n1.err; echo "eval exit $?"; grep "ms/frame" gpurun_out/eval_s17_n1.err | head -4; cat gpurun_out/eval_s17_n1.json
timeout 600 python tools/bench_hier_step.py --k 21 --iters 10 > gpurun_out/hier_step_s17.json 2> gpurun_out/hier_step_s17.err; echo "hier exit $?"; python -c "
import json;d=json.load(open('gpurun_out/hier_step_s17.json'));print({t:(e['ours'],e['bit_exact'],e['grad_rel_l2']) for t,e in d['taus'].items()})"
python __graft_entry__.py smoke 2>&1 | tail -1
