#!/bin/bash
# tools/quick.sh TAG: GPU tests + both bench workloads with per-kernel times -> gpurun_out/TAG_*
t=$1
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 --per-kernel --no-cpu > gpurun_out/${t}_bench.json 2> gpurun_out/${t}_bench.err
python bench.py --steps 10 --warmup 5 --per-kernel --no-cpu --workload 4k10qm > gpurun_out/${t}_bench_4k.json 2> gpurun_out/${t}_bench_4k.err
grep -h per-kernel gpurun_out/${t}_bench.err gpurun_out/${t}_bench_4k.err
python -c "
import json
for f in ('gpurun_out/${t}_bench.json','gpurun_out/${t}_bench_4k.json'):
    d=json.load(open(f)); print(round(d['value']), round(d['ms_per_step'],4), round(d['roofline']['frac'],3), round(d['roofline']['step_frac'],3), round(d['e2e']['value']))
"
