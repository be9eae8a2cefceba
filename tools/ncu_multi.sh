#!/bin/bash
# usage: tools/ncu_multi.sh <tag> <configs>   (run under gpurun; writes gpurun_out/<tag>.ncu-rep)
# configs = op:dtype:P:mode[,...]; full-set captures of the two launches of every configuration (NCU_SOURCE=1: with source)
# keep the list short: gpurun brings back at most 64 MiB
tag=$1; cfg=$2
python tools/quick_bench.py --configs "$cfg" --fixed --reps 1 --warmup 1 > gpurun_out/${tag}_plain.log 2>&1 || { tail -5 gpurun_out/${tag}_plain.log; exit 1; }
ncu --set full --clock-control none ${NCU_SOURCE:+--import-source on} -k regex:'column_kernel|column2_kernel|tiled_kernel|tc_kernel|cell_kernel' \
    -f -o gpurun_out/${tag} python tools/quick_bench.py --configs "$cfg" --fixed --reps 1 --warmup 1 > gpurun_out/${tag}_ncu.log 2>&1
tail -3 gpurun_out/${tag}_ncu.log
