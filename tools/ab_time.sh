#!/bin/bash
# plain timing A/B: [AB_ARGS="claims turns"] tools/ab_time.sh n_sessions reps lib1 [lib2 ...]
n=$1; r=$2; shift; shift
for lib in "$@"; do
  if [ "$lib" = default ]; then unset DVB_LIB; else export DVB_LIB=$PWD/$lib; fi
  echo "== $lib"; timeout 200 python tools/prof_bind.py $n $r $AB_ARGS 2>&1 | grep "kernel ms" | tail -3
done
