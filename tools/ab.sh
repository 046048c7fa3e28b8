#!/bin/bash
# tools/ab.sh TAG VARIANT... : on the GPU box, for each variant library tools/variants/libsvtgpu_V.so
# ("base" = the in-tree library): headline bench with parity gate and per-kernel table.
tag=$1; shift
for v in "$@"; do
  lib=""; [ "$v" != base ] && lib=$PWD/tools/variants/libsvtgpu_$v.so
  SVTGPU_LIB=$lib python bench.py --no-extra --no-cpu --per-kernel > gpurun_out/${tag}_${v}.json 2> gpurun_out/${tag}_${v}.err
  echo "== $v rc=$?"
  grep -h per-kernel gpurun_out/${tag}_${v}.err | cut -c1-700
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${tag}_${v}.json")); r = d.get("roofline") or {}
    print("$v", "value", d.get("value"), "ms", d.get("ms_per_step"), "frac", r.get("frac"), "step", r.get("step_frac"), "e2e", (d.get("e2e") or {}).get("value"), "parity", (d.get("parity") or {}).get("ok"))
except Exception as e:
    print("$v: no line", e)
PY
done
