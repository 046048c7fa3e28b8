#!/bin/bash
# stream / order knobs of svtgpu_run, kernel-only step time
for st in ${STREAMS:-2 3 4}; do for od in ${ORDERS:-0 1 2}; do
  echo -n "streams=$st order=$od: "
  SVTGPU_SIDE_STREAMS=$st SVTGPU_ORDER=$od python bench.py --steps 30 --warmup 5 --kernel-only 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(round(d['ms_per_step'],4), round(d['value']))"
done; done
