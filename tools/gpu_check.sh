#!/bin/bash
# tools/gpu_check.sh <tag> : on the GPU box -- the GPU parity tests, then the headline bench with the
# per-kernel table, then 4k10; everything lands in gpurun_out/<tag>_*
tag=${1:-run}
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/${tag}_tests.log
python bench.py --no-extra --no-cpu --per-kernel > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
grep per-kernel gpurun_out/${tag}_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_bench.json"))
r = d.get("roofline") or {}
print("1080p8", d.get("value"), "ms", d.get("ms_per_step"), "frac", r.get("frac"), "step", r.get("step_frac"), "launch_ms", r.get("launch_ms"), "e2e", (d.get("e2e") or {}).get("value"), d.get("parity"))
PY
python bench.py --no-extra --no-cpu --workload 4k10 --per-kernel > gpurun_out/${tag}_bench4k.json 2> gpurun_out/${tag}_bench4k.err
grep per-kernel gpurun_out/${tag}_bench4k.err
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_bench4k.json"))
r = d.get("roofline") or {}
print("4k10", d.get("value"), "ms", d.get("ms_per_step"), "frac", r.get("frac"), "step", r.get("step_frac"), d.get("parity"))
PY
