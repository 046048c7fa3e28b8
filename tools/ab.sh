#!/bin/bash
# tools/ab.sh OUTPREFIX VARIANT...   A/B the library builds under tools/variants on the GPU box
out=$1; shift
for v in "$@"; do
  export SVTGPU_LIB=$PWD/tools/variants/libsvtgpu_$v.so
  python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -2 > gpurun_out/${out}_${v}_pytest.log
  python bench.py --steps 20 --warmup 5 --per-kernel --no-cpu > gpurun_out/${out}_${v}.json 2> gpurun_out/${out}_${v}.err
  python bench.py --steps 20 --warmup 5 --kernel-only >> gpurun_out/${out}_${v}.json 2>> gpurun_out/${out}_${v}.err
  python bench.py --steps 10 --warmup 5 --per-kernel --no-cpu --workload 4k10qm > gpurun_out/${out}_${v}_4k.json 2> gpurun_out/${out}_${v}_4k.err
done
tail -n 3 gpurun_out/${out}_*_pytest.log
