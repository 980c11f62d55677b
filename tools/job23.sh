mkdir -p gpurun_out
for i in 1 2; do timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-cut > gpurun_out/b$i.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/b$i.json')); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'])"; done
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bref.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/bref.json')); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])"
nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.width.current --format=csv
