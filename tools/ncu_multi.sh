#!/bin/bash
# usage: tools/ncu_multi.sh <tag> <configs>   (run under gpurun; writes gpurun_out/<tag>.ncu-rep)
# configs = op:dtype:P:mode[,...]; one full-set capture of the last launch of every configuration
tag=$1; cfg=$2
python tools/quick_bench.py --configs "$cfg" --reps 1 --warmup 1 > gpurun_out/${tag}_plain.log 2>&1 || { tail -5 gpurun_out/${tag}_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'column_kernel|tiled_kernel|tc_kernel' \
    -f -o gpurun_out/${tag} python tools/quick_bench.py --configs "$cfg" --reps 1 --warmup 0 > gpurun_out/${tag}_ncu.log 2>&1
tail -3 gpurun_out/${tag}_ncu.log
