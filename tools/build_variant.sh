#!/bin/bash
# tools/build_variant.sh NAME [extra nvcc flags...]  ->  tools/variants/libsvtgpu_NAME.so
# Experiment builds of the library (A/B runs on the GPU box: SVTGPU_LIB=<path> python bench.py ...).
set -e
cd "$(dirname "$0")/.."
name=$1; shift
out=tools/variants; obj=$out/obj_$name
mkdir -p $obj
pids=()
for f in cidana_svt_av1_b200/csrc/*.cu; do
  /usr/local/cuda/bin/nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC "$@" \
     -c $f -o $obj/$(basename ${f%.cu}).o &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
/usr/local/cuda/bin/nvcc -shared -o $out/libsvtgpu_$name.so $obj/*.o
rm -rf $obj
echo built $out/libsvtgpu_$name.so
