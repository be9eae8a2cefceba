#!/bin/bash
# usage: tools/ncu_one.sh <tag> <quick_bench args...>   (run under gpurun; writes gpurun_out/<tag>.ncu-rep)
tag=$1; shift
python tools/quick_bench.py "$@" --reps 2 > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'column_kernel|tiled_kernel' -s 3 -c 2 \
    -f -o gpurun_out/${tag} python tools/quick_bench.py "$@" --reps 2 > gpurun_out/${tag}_ncu.log 2>&1
tail -2 gpurun_out/${tag}_plain.log
