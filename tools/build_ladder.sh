#!/bin/bash
# Cost-ladder builds of the bind kernel: _ab/rung<n>.so = libdvb200 with bind.cu compiled -DB7_RUNG=n (bind7.cuh).
# Only k_bind_facts differs (rungs stop its fact walk early); rung 6 is the product.  Run tools/ladder_run.sh on the GPU box afterwards.
set -e
cd "$(dirname "$0")/.."
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --fmad=true -Xcompiler -fPIC -Xcompiler -O2"
mkdir -p _ab/obj
for f in deep_vision_b200/csrc/*.cu; do
  b=$(basename $f .cu); [ $b = bind ] && continue
  nvcc $FLAGS -c $f -o _ab/obj/$b.o &
done
wait
for n in ${@:-0 1 2 3 4 5 6}; do
  ( nvcc $FLAGS -DB7_RUNG=$n -c deep_vision_b200/csrc/bind.cu -o _ab/obj/bind_r$n.o &&
    nvcc -shared -cudart static -gencode arch=compute_100a,code=sm_100a -o _ab/rung$n.so _ab/obj/bind_r$n.o $(ls _ab/obj/*.o | grep -v bind_r) ) &
done
wait
ls -la _ab/rung*.so
