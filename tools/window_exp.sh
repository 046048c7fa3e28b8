#!/bin/bash
for w in 10 12 13 14 16 20; do
  echo -n "window_log2=$w: "
  SVTGPU_SORT_WINDOW_LOG2=$w python bench.py --steps 30 --warmup 5 --kernel-only 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(round(d['ms_per_step'],4), round(d['value']))"
done
