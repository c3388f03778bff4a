#!/bin/bash
# A/B builds of the bind kernel: tools/build_variants.sh name1:"-DX=0 -DY=1" name2:"..." -> _ab/<name>.so (other objects shared)
set -e
cd "$(dirname "$0")/.."
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --fmad=true -Xcompiler -fPIC -Xcompiler -O2"
mkdir -p _ab/obj
for f in deep_vision_b200/csrc/*.cu; do
  b=$(basename $f .cu); [ $b = bind ] && continue
  [ _ab/obj/$b.o -nt $f ] || nvcc $FLAGS -c $f -o _ab/obj/$b.o &
done
wait
for spec in "$@"; do
  name=${spec%%:*}; defs=${spec#*:}; [ "$defs" = "$spec" ] && defs=""
  ( nvcc $FLAGS $defs -Xptxas -v -c deep_vision_b200/csrc/bind.cu -o _ab/obj/bind_$name.o 2>&1 | grep -A2 "k_bind_claims10" | grep -E "spill|Used" | tr '\n' ' ' | sed "s/^/$name: /"; echo
    nvcc -shared -cudart static -gencode arch=compute_100a,code=sm_100a -o _ab/$name.so _ab/obj/bind_$name.o $(ls _ab/obj/*.o | grep -v "/bind_") ) &
done
wait
