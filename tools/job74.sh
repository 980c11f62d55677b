set -x
mkdir -p gpurun_out
for k in 24 21; do
timeout 600 python tools/bench_cut.py --k $k --frames 30 > gpurun_out/cut_k${k}_batched.json 2> gpurun_out/cut_k${k}_batched.err; echo "exit $?"
LODGS_CUT_FLAT_ROWWISE=1 timeout 600 python tools/bench_cut.py --k $k --frames 30 > gpurun_out/cut_k${k}_rowwise.json 2> gpurun_out/cut_k${k}_rowwise.err; echo "exit $?"
done
python - <<'PY'
import json
for k in (24, 21):
    for v in ("batched", "rowwise"):
        d = json.load(open("gpurun_out/cut_k%d_%s.json" % (k, v)))
        print(k, v, {t: (e["flat"]["cut_ms_per_frame"], e["traversal"]["cut_ms_per_frame"], e["bit_exact_vs_reference"]) for t, e in d["taus"].items()})
PY
