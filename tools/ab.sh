#!/bin/bash
# usage: tools/ab.sh <out> "<quick_bench args>" <suffix> [<suffix>...]   ("" = product library)
out=$1; args=$2; shift 2
: > $out
for sfx in "$@"; do
  echo "== lib${sfx:-(product)}" >> $out
  if [ -z "$sfx" ] || [ "$sfx" = "-" ]; then python tools/quick_bench.py $args >> $out 2>&1
  else FEMTC_B200_LIB=$PWD/femtc_b200/libfemtc_b200${sfx}.so python tools/quick_bench.py $args >> $out 2>&1; fi
done
